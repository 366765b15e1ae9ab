#!/usr/bin/env python
"""bench.py — the multi-source-BFS hot path on synthetic BASELINE.json workloads.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

A step = one pass of the hot path over the whole query set of the workload: every group's
BFS + distance sum (what the reference does in its "Computation time" window,
/root/reference/main.cu:301-399).  With N > 1 (one process per GPU, torchrun) the groups
are sharded in contiguous 512-lane batches over the ranks (strong scaling: total work
fixed), the graph is replicated (each rank generates it on its own GPU), and the only
exchange is one 8-byte packed min all-reduce per step.

Reported: metric GTEPS (sum over groups of traversed edges / time) with queries/s beside
it; `value` with queries and sums resident in HBM, device-timed by CUDA events (max over
ranks); `e2e` through msbfs_run with host buffers (H2D of the queries, D2H of the sums
inside the timed region); the roofline of the dominant kernel from live CUDA-event
timings; a CPU baseline (the oracle port, OpenMP over groups) on a bounded sample.
`--impl reference` times the reference's own CPU implementation: oracle/_ref/main_host
(the unmodified main.cu built for the host) under the bundled mpirun where it can load
the workload, else the oracle port.  Only the cpu_baseline / reference legs touch oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

WORKLOADS = {
    # name: graph generator + query shape (BASELINE.json configs 1-5)
    "rmat16": dict(kind="rmat", p0=16, p1=0, m=1 << 20, groups=64, per_group=4,
                   desc="RMAT scale-16 ef16, 64 groups x 4 sources (config 1)"),
    "rmat22": dict(kind="rmat", p0=22, p1=0, m=1 << 26, groups=1024, per_group=8,
                   desc="RMAT scale-22 ef16, 1024 groups x 8 sources (config 2)"),
    "kron26": dict(kind="kron", p0=26, p1=0, m=(1 << 30) - 1, groups=4096, per_group=16,
                   desc="Graph500 Kronecker scale-26 ef16 (2^30-1 records), 4096 groups x 16 sources (config 3)"),
    "lattice4096": dict(kind="lattice", p0=4096, p1=4096, m=2 * 4096 * 4095, groups=512, per_group=4,
                        desc="2D lattice 4096x4096, 512 groups x 4 sources (config 4)"),
    "er24": dict(kind="er", p0=1 << 24, p1=0, m=1 << 28, groups=1024, per_group=256,
                 desc="Erdos-Renyi n=2^24 m=2^28, 1024 groups x 256 sources (config 5)"),
    # reduced shapes for quick checks (not BASELINE configs)
    "kron22": dict(kind="kron", p0=22, p1=0, m=(1 << 26), groups=1024, per_group=16,
                   desc="Kronecker scale-22 ef16, 1024 groups x 16 sources (reduced, not a BASELINE config)"),
    "lattice1024": dict(kind="lattice", p0=1024, p1=1024, m=2 * 1024 * 1023, groups=256, per_group=4,
                        desc="2D lattice 1024x1024, 256 groups x 4 sources (reduced, not a BASELINE config)"),
}
GRAPH_SEED, QUERY_SEED = 1, 2
KINDS = {"rmat": 0, "kron": 1, "lattice": 2, "er": 3}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# The driver reads ONE JSON line from stdout.  Libraries (NCCL's version banner, torchrun
# notices) may print to fd 1, so fd 1 is pointed at stderr for the whole run and the result
# line goes to a private duplicate of the original stdout.
_RESULT_FD = None


def claim_stdout():
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); smax.append(float(r[1]))
            except Exception:
                continue
            for name, val in zip(names, r[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


# ------------------------------------------------------------------------ reference arm

def oracle_csr(n, uv_host):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    ob.set_threads(ob.host_threads())  # launchers export OMP_NUM_THREADS=1; state the team size ourselves
    t0 = time.time()
    csr = ob.Csr(n, uv_host, parallel=uv_host.shape[0] > (1 << 24))
    log(f"[cpu] oracle CSR built in {time.time() - t0:.1f} s on {ob.host_threads()} threads")
    return csr


def oracle_sample_rate(pkg, csr, offs, src, threads, target_seconds, pick_order=None):
    """Time the oracle port (BFS + sum only, CSR prebuilt) on a bounded sample of groups.
    `pick_order`: group ids to draw the sample from, in that order (default: file order).
    Returns (sums, reached_entries, seconds, group ids timed, threads)."""
    import oracle_binding as ob

    k_total = offs.size - 1
    threads = threads or ob.host_threads()
    order = np.arange(k_total) if pick_order is None else np.asarray(pick_order)
    # probe with one group per thread, then size the sample for ~target_seconds
    probe = min(k_total, threads)

    def timed(ids):
        o, s_ = pkg.gather_groups(offs, src, ids)
        t0 = time.time()
        r = ob.run_csr(csr, o, s_, threads)
        return r, max(time.time() - t0, 1e-6)

    res, t_probe = timed(order[:probe])
    rounds = max(1, int(target_seconds / t_probe))
    sample = min(k_total, probe * rounds)
    if sample > probe:
        res, t_probe = timed(order[:sample])
    return res.sums, res.reached_entries, t_probe, order[:sample], threads


def spread_over_run_order(costs, lanes=512):
    """Group ids interleaved from the first, a middle and the last batch of the engine's run
    order (heaviest first): a sample cut from the head of this list checks the hub-source
    batch, a middle one and the far-from-hub tail, not just the head of the file."""
    order = np.lexsort((np.arange(costs.size), -np.asarray(costs, dtype=np.int64)))
    nb = (order.size + lanes - 1) // lanes
    parts = [order[b * lanes:(b + 1) * lanes] for b in sorted({0, nb // 2, nb - 1})]
    out = []
    for i in range(max(len(p_) for p_ in parts)):
        for p_ in parts:
            if i < len(p_):
                out.append(int(p_[i]))
    seen = set(out)
    rest = [int(g) for g in order if int(g) not in seen]
    return np.array(out + rest, dtype=np.int64)


def run_reference_arm(args, wl, rank):
    """bench.py --impl reference: the reference's CPU implementation of the path on this
    box's host cores, same workload / metric / unit as our arm."""
    if rank != 0:
        return
    pkg = graft.load_package()
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    kind = KINDS[wl["kind"]]
    n = pkg.gen_num_vertices(kind, wl["p0"], wl["p1"])
    cores = ob.host_threads()
    ob.set_threads(cores)
    t0 = time.time()
    uv = make_edges_host(pkg, wl, kind)
    log(f"[ref] edges ready in {time.time() - t0:.1f} s")
    offs, src = pkg.gen_queries_host(n, wl["groups"], wl["per_group"], QUERY_SEED)
    use_ref_binary = os.path.exists(ob.REF_HOST) and wl["m"] <= (1 << 26) and wl["per_group"] <= 255 and not args.force_port
    values, qps, step_s = [], [], []
    if use_ref_binary:
        # the unmodified main.cu (host build) under the bundled mpirun, <= 255 groups per file
        ranks = max(1, min(cores, 16))
        sample = min(wl["groups"], 255, max(ranks, args.ref_groups))
        csr = ob.Csr(n, uv, parallel=True)
        reached = ob.run_csr(csr, offs[: sample + 1], src, cores).reached_entries
        with tempfile.TemporaryDirectory() as tmp:
            g, q = os.path.join(tmp, "g.bin"), os.path.join(tmp, "q.bin")
            pkg.write_graph_file(g, n, uv)
            pkg.write_query_file(q, offs[: sample + 1], src[: offs[sample]])
            mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
            for it in range(args.warmup + args.steps):
                res = subprocess.run([mpirun, "-np", str(ranks), ob.REF_HOST, "-g", g, "-q", q, "-gn", "1"],
                                     capture_output=True, text=True)
                if res.returncode != 0:
                    raise RuntimeError(res.stderr)
                comp = float([ln for ln in res.stdout.splitlines() if ln.startswith("Computation time")][0].split()[2])
                if it >= args.warmup:
                    values.append(reached.sum() / 2 / comp / 1e9)
                    qps.append(sample / comp)
                    step_s.append(comp)
        kind_s, cores_used = "reference", ranks
        sample_desc = (f"first {sample} of {wl['groups']} groups, oracle/_ref/main_host (unmodified main.cu, host build) "
                       f"under mpirun -np {ranks}; its own 'Computation time' line")
    else:
        # The oracle port with all host threads.  One BFS per thread is the faster layout, but a
        # round of it must fit the per-step budget (a Kronecker-26 BFS takes a thread ~25 s); when
        # it does not, groups run one after the other with every thread inside each BFS.
        threads = cores
        csr = oracle_csr(n, uv)
        del uv
        iters = args.warmup + args.steps
        budget = max(1.0, args.ref_seconds / iters)
        o1, s1 = pkg.gather_groups(offs, src, np.arange(1))
        t0 = time.time()
        ob.run_csr(csr, o1, s1, threads, wide=True)
        t_wide = max(time.time() - t0, 1e-6)
        wide = t_wide * min(threads, 8) > budget   # a round of `threads` single-thread BFS would not fit
        per_step = max(1, int(budget / t_wide)) if wide else threads * max(1, int(budget / max(t_wide * min(threads, 8), 1e-6)))
        per_step = min(per_step, wl["groups"])
        log(f"[ref] one BFS on {threads} threads: {t_wide:.2f} s; {per_step} groups per step, "
            f"{'all threads inside each BFS' if wide else 'one BFS per thread'}")
        for it in range(iters):
            ids = (np.arange(per_step) + it * per_step) % wl["groups"]
            o_, s_ = pkg.gather_groups(offs, src, ids)
            t0 = time.time()
            res = ob.run_csr(csr, o_, s_, threads, wide=wide)
            secs = max(time.time() - t0, 1e-6)
            if it >= args.warmup:
                values.append(res.reached_entries.sum() / 2 / secs / 1e9)
                qps.append(per_step / secs)
                step_s.append(secs)
        kind_s, cores_used = "port", threads
        sample = per_step
        sample_desc = (f"{per_step} of {wl['groups']} groups per step (consecutive groups of the file, a new set each step), "
                       f"oracle port (oracle/oracle.c), {threads} threads, "
                       + ("all threads inside each BFS; " if wide else "one BFS per thread; ")
                       + "the host build of main.cu cannot load this workload in bounded time")
    v = statistics.median(values)
    line = {
        "impl": "reference", "metric": "GTEPS", "value": v, "unit": "GTEPS", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": statistics.median(step_s) * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic",
        "config": {"workload": args.workload, "description": wl["desc"], "graph_seed": GRAPH_SEED, "query_seed": QUERY_SEED},
        "queries_per_sec": statistics.median(qps),
        "cpu_baseline": {"value": v, "unit": "GTEPS", "cores": cores_used, "kind": kind_s, "sample": sample_desc},
        "e2e": {"value": v, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def make_edges_host(pkg, wl, kind):
    """Host copy of the workload's edge records: from the GPU generator when there is a GPU
    (seconds even for 8 GiB), else from the identical CPU generator."""
    try:
        import torch
        if torch.cuda.is_available():
            buf = torch.empty((wl["m"], 2), dtype=torch.int32, device="cuda:0")
            rc = pkg.lib().msbfs_gen_edges_device(0, kind, wl["p0"], wl["p1"], wl["m"], GRAPH_SEED, buf.data_ptr())
            if rc == 0:
                out = buf.cpu().numpy()
                del buf
                torch.cuda.empty_cache()
                return out
    except Exception as e:  # pragma: no cover
        log(f"[ref] device generator unavailable ({e}); generating on the host")
    return pkg.gen_edges_host(kind, wl["p0"], wl["p1"], wl["m"], GRAPH_SEED)


def reference_gpu_block(pkg, args, wl, n, uv_host, our_qps):
    """The reference's own GPU program — unmodified main.cu built for sm_100a (oracle/_ref/main_gpu)
    — on this box, on the same graph file and the first <= 255 groups (its query format's limit,
    main.cu:143-152); its own "Computation time" line (main.cu:412-413).  A reported baseline."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    if not os.path.exists(ob.REF_GPU):
        return {"unavailable": "oracle/_ref/main_gpu not built (needs /root/reference at build time)"}
    if (wl["m"] > (1 << 26) or wl["per_group"] > 255) and not args.ref_gpu_force:
        return {"unavailable": "the reference's loader (one fread + push_back per id, main.cu:108-116) needs minutes for "
                               f"{wl['m']} records" + (" and its query format caps a group at 255 sources" if wl["per_group"] > 255 else "")
                               + "; run with --ref-gpu-force to wait for it"}
    groups = min(args.ref_groups, wl["groups"], 255)
    per_group = min(wl["per_group"], 255)
    offs, src = pkg.gen_queries_host(n, wl["groups"], wl["per_group"], QUERY_SEED)
    sel = [src[offs[i]:offs[i] + per_group] for i in range(groups)]
    offs2 = np.concatenate([[0], np.cumsum([len(x) for x in sel])]).astype(np.int64)
    src2 = np.concatenate(sel).astype(np.int32)
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as tmp:
        g, q = os.path.join(tmp, "graph.bin"), os.path.join(tmp, "query.bin")
        pkg.write_graph_file(g, n, uv_host)
        pkg.write_query_file(q, offs2, src2)
        mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
        out = {}
        for name, exe in (("reference", ob.REF_GPU), ("drop_in", os.path.join(pkg.PKG_DIR, "host", "main"))):
            try:
                res = subprocess.run([mpirun, "-np", "1", exe, "-g", g, "-q", q, "-gn", "1"], capture_output=True, text=True,
                                     timeout=args.ref_gpu_timeout)
            except subprocess.TimeoutExpired:
                return {"unavailable": f"{name} did not finish within {args.ref_gpu_timeout:.0f} s"}
            if res.returncode != 0:
                return {"unavailable": f"{name} failed: {res.stderr[-200:]}"}
            rec = {}
            for ln in res.stdout.splitlines():
                if ln.startswith("Query number"): rec["k_printed"] = int(ln.split(":")[1])
                elif ln.startswith("Minimum F value"): rec["min_f"] = int(ln.split(":")[1])
                elif ln.startswith("Preprocessing time"): rec["preprocessing_s"] = float(ln.split()[2])
                elif ln.startswith("Computation time"): rec["computation_s"] = float(ln.split()[2])
            out[name] = rec
    r, d = out["reference"], out["drop_in"]
    return {"program": "oracle/_ref/main_gpu = unmodified /root/reference/main.cu, nvcc sm_100a, mpirun -np 1 -gn 1",
            "sample": f"first {groups} groups x {per_group} sources of the workload (reference query format)",
            "reference": r, "drop_in_main_same_files": d,
            "queries_per_sec": groups / r["computation_s"] if r.get("computation_s") else None,
            "same_answer": (r.get("k_printed"), r.get("min_f")) == (d.get("k_printed"), d.get("min_f")),
            "ours_over_reference_queries_per_sec": (our_qps / (groups / r["computation_s"])) if r.get("computation_s") else None}


# ------------------------------------------------------------------------------ our arm

def run_ours(args, wl, rank, local_rank, world):
    import torch
    import torch.distributed as dist

    pkg = graft.load_package()
    import importlib
    sharding = importlib.import_module(pkg.__name__ + ".sharding")

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    kind = KINDS[wl["kind"]]
    n = pkg.gen_num_vertices(kind, wl["p0"], wl["p1"])
    m = wl["m"]
    K = wl["groups"]
    offs, src = pkg.gen_queries_host(n, K, wl["per_group"], QUERY_SEED)

    # ---- graph: generated on this rank's GPU (replicated), CSR built on the device
    t0 = time.time()
    buf = torch.empty((m, 2), dtype=torch.int32, device=dev)
    rc = pkg.lib().msbfs_gen_edges_device(local_rank, kind, wl["p0"], wl["p1"], m, GRAPH_SEED, buf.data_ptr())
    if rc != 0:
        raise RuntimeError(pkg.lib().msbfs_last_error().decode())
    gen_s = time.time() - t0
    eng = pkg.Engine(local_rank)
    eng.set_option("profile", 1)
    if args.alpha:
        eng.set_option("alpha", args.alpha)
    for kv in args.opt:
        key, val = kv.split("=")
        eng.set_option(key, int(val))
    if args.lanes:
        eng.set_option("lanes", args.lanes)
    t0 = time.time()
    eng.set_graph_edges_device(n, m, buf.data_ptr())
    build_s = time.time() - t0
    want_cpu = (rank == 0 and world == 1 and not args.no_cpu)
    uv_host = None
    if want_cpu:
        t0 = time.time()
        uv_host = buf.cpu().numpy()
        log(f"[cpu] edge list copied to the host in {time.time() - t0:.1f} s")
    del buf
    torch.cuda.empty_cache()

    # Groups are ordered by estimated BFS "speed" (msbfs_group_weights, heaviest first)
    # and every rank takes a contiguous run of that order: lanes that reach the hubs at the same
    # level finish their levels together, so early exit works for the whole batch.  Kronecker-26:
    # scanned entries per step 19.5 G -> 14.9 G, lane words fetched 7.9 G -> 4.3 G.
    costs = eng.group_weights(offs, src)
    if args.contiguous:
        lo_, hi_ = pkg.shard_range(K, rank, world)
        mine = np.arange(lo_, hi_)
    else:
        # batches of that order are handed to the ranks in snake order (fast and slow batches mixed)
        share, plan_lanes = pkg.plan_shares(costs, world)
        mine = share[rank].astype(np.int64)
        if plan_lanes and not args.lanes:
            eng.set_option("lanes", plan_lanes)
    my_offs, my_src = pkg.gather_groups(offs, src, mine)
    lo, hi = 0, int(mine.size)   # local group range
    h2d_bytes = int(my_offs.nbytes + my_src.nbytes)
    d2h_bytes = int(mine.size * 8)
    log(f"[rank {rank}] n={n} m={m} {mine.size} groups gen {gen_s:.2f}s build {build_s:.2f}s "
        f"(device {eng.stats()['build_ms']:.0f} ms)")

    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def reduce_max(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def reduce_sum(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- traversed edges per group (TEPS numerator), one untimed pass
    eng.upload_queries(my_offs, my_src)
    reached = eng.reached_entries() if hi > lo else np.zeros(0, np.int64)
    traversed_edges_total = reduce_sum(float(reached.sum()) / 2.0)

    def one_step_resident():
        """queries and sums resident in HBM; returns device ms (CUDA events in the library)"""
        if hi > lo:
            eng.run_resident()
            ms = eng.stats()["run_ms"]
            sums = eng.download_sums()
        else:
            ms, sums = 0.0, np.zeros(0, np.int64)
        packed = pkg.argmin_packed_ids(sums, mine)
        best = sharding.allreduce_packed_min(packed, dev)
        return ms, sums, best

    # ---- warm-up
    for _ in range(args.warmup):
        flush_buf.zero_()
        one_step_resident()

    # ---- plan from measured cost: the warm-up's per-batch device times (every rank's, in global
    #      batch order) go back into msbfs_plan_shares, which hands the batches out longest first
    #      to the least-loaded rank instead of in snake order
    plan_note = "one rank" if world == 1 else ("contiguous file order" if args.contiguous else "snake order")

    def adopt(new_share, lanes_opt):
        nonlocal share, mine, my_offs, my_src, lo, hi, h2d_bytes, d2h_bytes
        share = new_share
        mine = share[rank].astype(np.int64)
        my_offs, my_src = pkg.gather_groups(offs, src, mine)
        lo, hi = 0, int(mine.size)
        h2d_bytes = int(my_offs.nbytes + my_src.nbytes)
        d2h_bytes = int(mine.size * 8)
        if lanes_opt:
            eng.set_option("lanes", lanes_opt)
        eng.upload_queries(my_offs, my_src)

    def warm_max_ms(n=2):
        out = 0.0
        for _ in range(n):
            flush_buf.zero_()
            out = reduce_max(one_step_resident()[0])
        return out

    # ---- batch width by measurement: when 512-group batches would leave every rank exactly one, the
    #      static rule picks 256 on skewed weights (a rank then mixes a fast and a slow batch).  Whether
    #      that pays depends on the graph (Kronecker-26 at 8 GPUs: yes, by 10 %; RMAT-22 at 2: no, by
    #      30 %), so the other width gets two warm-up steps too and the faster plan is kept.
    width_note = ""
    if (world > 1 and not args.contiguous and not args.no_replan and not args.lanes and plan_lanes == 256
            and (K + 511) // 512 >= world):
        t256 = warm_max_ms(1)
        share256 = share
        share512, _ = pkg.plan_shares(costs, world, lanes=512)
        adopt(share512, 512)
        t512 = warm_max_ms(2)
        if t512 <= t256:
            plan_lanes = 512
            width_note = f"; 512-lane batches measured faster than 256 ({t512:.1f} vs {t256:.1f} ms)"
        else:
            adopt(share256, 256)
            warm_max_ms(1)
            width_note = f"; 256-lane batches measured faster than 512 ({t256:.1f} vs {t512:.1f} ms)"

    if world > 1 and not args.contiguous and not args.no_replan:
        mine_ms = eng.batch_ms() if hi > lo else []
        all_ms = [None] * world
        dist.all_gather_object(all_ms, mine_ms)
        width = plan_lanes or 512
        nb = (K + width - 1) // width
        order_all = np.concatenate(share)  # not in run order; rebuild it from the plan
        run_order = np.lexsort((np.arange(K), -np.asarray(costs, dtype=np.int64)))
        owner_of_batch = []
        owner = np.empty(K, np.int64)
        for r in range(world):
            owner[share[r]] = r
        cost_b, seen = [], [0] * world
        for b in range(nb):
            r = int(owner[run_order[b * width]])
            owner_of_batch.append(r)
            cost_b.append(float(all_ms[r][seen[r]]) if seen[r] < len(all_ms[r]) else 0.0)
            seen[r] += 1
        share2, _ = pkg.plan_shares(costs, world, lanes=width, batch_cost=np.array(cost_b))
        loads_old = [sum(c for c, o in zip(cost_b, owner_of_batch) if o == r) for r in range(world)]
        owner2 = np.empty(K, np.int64)
        for r in range(world):
            owner2[share2[r]] = r
        loads_new = [sum(c for b, c in enumerate(cost_b) if int(owner2[run_order[b * width]]) == r) for r in range(world)]
        if max(loads_new) < 0.98 * max(loads_old):
            adopt(share2, 0)
            flush_buf.zero_()
            one_step_resident()  # the new share's own warm-up
            plan_note = (f"longest batch first onto the least-loaded rank, from the warm-up's per-batch times "
                         f"(planned max per rank {max(loads_new):.1f} ms vs {max(loads_old):.1f} ms in snake order)")
        else:
            plan_note = f"snake order (measured-cost plan no better: {max(loads_new):.1f} vs {max(loads_old):.1f} ms)"
        del order_all
        plan_note += width_note

    # ---- timed steps: device time (events) per step, max over ranks; wall as a cross-check
    sampler = ClockSampler(local_rank)
    sync_all()
    sampler.start()
    step_ms, prof_acc, launches = [], None, 0
    wall0 = time.time()
    for _ in range(args.steps):
        flush_buf.zero_()
        torch.cuda.synchronize(dev)
        ms, sums, best = one_step_resident()
        step_ms.append(ms)
        launches += eng.stats()["kernel_launches"] if hi > lo else 0
        prof = eng.kernel_profiles() if hi > lo else []
        if prof_acc is None:
            prof_acc = prof
        else:
            for a_, b_ in zip(prof_acc, prof):
                for key in ("launches", "ms", "alg_bytes", "scanned", "words_fetched", "vertices_examined", "words_written"):
                    a_[key] += b_[key]
    sync_all()
    wall_s = time.time() - wall0
    clocks = sampler.stop()
    device_ms = reduce_max(sum(step_ms))
    ms_per_step = device_ms / args.steps
    per_rank_ms = [sum(step_ms) / args.steps]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, per_rank_ms[0])
        per_rank_ms = [float(x) for x in gathered]
    batch_ms_rank0 = [round(x, 3) for x in eng.batch_ms()] if hi > lo else []
    stats = eng.stats() if hi > lo else {}
    min_k, min_f = pkg.unpack_argmin(best)

    # ---- e2e: the public call with host buffers (H2D queries + D2H sums inside the timing)
    sync_all()
    e2e_t = []
    for _ in range(args.steps):
        flush_buf.zero_()
        sync_all()
        t0 = time.perf_counter()
        sums_e = eng.run(my_offs, my_src) if hi > lo else np.zeros(0, np.int64)
        best_e = sharding.allreduce_packed_min(pkg.argmin_packed_ids(sums_e, mine), dev)
        torch.cuda.synchronize(dev)
        e2e_t.append(time.perf_counter() - t0)
    e2e_s = reduce_max(sum(e2e_t)) / args.steps
    assert best_e == best

    if rank == 0 and args.level_log:
        with open(args.level_log, "w") as f:
            for rec in eng.level_log():
                f.write(json.dumps(rec) + "\n")

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (live CUDA-event timings of the timed steps)
    peak, peak_src = measured_peaks()
    roofline = None
    kernels = []
    if prof_acc:
        for p in prof_acc:
            if p["launches"]:
                kernels.append({"name": p["name"], "launches": p["launches"], "ms": round(p["ms"], 3),
                                "alg_gb": round(p["alg_bytes"] / 1e9, 3),
                                "alg_gbs": round(p["alg_bytes"] / 1e6 / p["ms"], 1) if p["ms"] > 0 else None})
        top = max(prof_acc, key=lambda p: p["ms"])
        if top["ms"] > 0 and top["launches"] > 0:
            achieved = top["alg_bytes"] / 1e9 / (top["ms"] / 1e3)
            traffic = None
            tpath = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tpath):
                try:
                    traffic = json.load(open(tpath)).get(args.workload, {}).get(top["name"])
                except Exception:
                    traffic = None
            avg_ms = top["ms"] / top["launches"]
            roofline = {"bound": "hbm", "kernel": top["name"], "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                        "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                        "frac_note": "achieved = ALGORITHMIC bytes / time (SURVEY 8(d)); dram_frac = ncu DRAM bytes per launch "
                                     "(traffic, profiles/traffic.json) / the same time: hub words are served from L2",
                        "dram_frac": round(traffic / 1e9 / (avg_ms / 1e3) / peak, 4) if traffic else None,
                        "alg_bytes_per_launch": int(top["alg_bytes"] / top["launches"]),
                        "avg_launch_ms": round(top["ms"] / top["launches"], 4),
                        "whole_step": {"alg_bytes": int(stats.get("alg_bytes", 0)),
                                       "achieved": round(stats.get("alg_bytes", 0) / 1e9 / (ms_per_step / 1e3), 1) if ms_per_step else None,
                                       "frac": round(stats.get("alg_bytes", 0) / 1e9 / (ms_per_step / 1e3) / peak, 4) if ms_per_step else None,
                                       "note": "rank 0's algorithmic bytes of one step / its device time"}}

    # ---- CPU baseline (oracle port) on a bounded sample, N = 1 only
    cpu = None
    parity = None
    if want_cpu:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_binding as ob_
        o_sums, o_reached, secs, ids, threads = oracle_sample_rate(pkg, oracle_csr(n, uv_host), offs, src,
                                                                    ob_.host_threads(), args.cpu_seconds,
                                                                    pick_order=spread_over_run_order(costs))
        sample = int(ids.size)
        cpu = {"value": float(o_reached.sum()) / 2 / secs / 1e9, "unit": "GTEPS", "cores": threads, "kind": "port",
               "queries_per_sec": sample / secs,
               "sample": f"{sample} of {K} groups taken in turn from the first, a middle and the last 512-group batch of the "
                         f"run order, oracle port (one BFS per thread), {secs:.1f} s of BFS+sum"}
        full_sums = np.zeros(K, np.int64); full_sums[mine] = sums       # back to the file's group order
        full_reached = np.zeros(K, np.int64); full_reached[mine] = reached
        ok = np.array_equal(o_sums, full_sums[ids]) and np.array_equal(o_reached, full_reached[ids])
        parity = {"groups_checked": sample, "bit_exact": bool(ok),
                  "groups": "first / middle / last batch of the run order, interleaved"}
        if not ok:
            log("PARITY FAILURE against the oracle on the sampled groups")

    ref_gpu = None
    if want_cpu:
        try:
            ref_gpu = reference_gpu_block(pkg, args, wl, n, uv_host, K / (ms_per_step / 1e3))
        except Exception as e:  # a reported baseline must not take the bench line down
            ref_gpu = {"unavailable": f"{type(e).__name__}: {e}"}
        if ref_gpu and ref_gpu.get("same_answer") is False:
            log("PARITY FAILURE: the drop-in and the reference's GPU build disagree on the same files")
            parity = dict(parity or {}, bit_exact=False)

    value = traversed_edges_total / (ms_per_step / 1e3) / 1e9
    line = {
        "metric": "GTEPS", "value": value, "unit": "GTEPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64", "data": "synthetic",
        "config": {"workload": args.workload, "description": wl["desc"], "graph_seed": GRAPH_SEED, "query_seed": QUERY_SEED,
                   "lanes": stats.get("lanes"), "batches_rank0": stats.get("batches"), "levels_rank0": stats.get("levels"),
                   "levels_top_down": stats.get("levels_top_down"), "levels_bottom_up": stats.get("levels_bottom_up"),
                   "parallelism": f"query groups ordered heaviest first, cut into batches, batches over {world} rank(s) ({plan_note}), graph replicated",
                   "l2": "flushed between steps (512 MiB write); CSR larger than L2" if m * 8 > (126 << 20) else "flushed between steps (512 MiB write)"},
        "queries_per_sec": K / (ms_per_step / 1e3),
        "traversed_edges_per_step": traversed_edges_total,
        "argmin": {"k_printed": min_k + 1, "min_f": min_f},
        "wall_ms_per_step": wall_s * 1e3 / args.steps,
        "preprocessing": {"generate_s": round(gen_s, 3), "csr_build_s": round(build_s, 3),
                          "csr_build_device_ms": round(eng.stats()["build_ms"], 1)},
        "e2e": {"value": traversed_edges_total / e2e_s / 1e9, "unit": "GTEPS", "h2d_bytes_per_step": h2d_bytes,
                "d2h_bytes_per_step": d2h_bytes, "queries_per_sec": K / e2e_s, "ms_per_step": e2e_s * 1e3},
        "per_rank_device_ms": [round(x, 3) for x in per_rank_ms],
        "batch_ms_rank0": batch_ms_rank0,
        "plan": plan_note,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "kernels": kernels,
        "cpu_baseline": cpu,
        "reference_gpu": ref_gpu,
        "parity_sample": parity,
    }
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity is not None and not parity["bit_exact"]:
        raise SystemExit(3)  # a fast wrong answer is not a result


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("MSBFS_BENCH_WORKLOAD", "kron26"), choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work per oracle timing (bounded sample)")
    ap.add_argument("--ref-groups", type=int, default=32, help="groups fed to the reference host build per step")
    ap.add_argument("--ref-seconds", type=float, default=200.0, help="reference arm: CPU seconds for all its steps together")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline and reference_gpu legs")
    ap.add_argument("--ref-gpu-force", action="store_true", help="run the reference's GPU build even where its loader takes minutes")
    ap.add_argument("--ref-gpu-timeout", type=float, default=1500.0)
    ap.add_argument("--force-port", action="store_true", help="reference arm: always use the oracle port")
    ap.add_argument("--level-log", default="", help="write the per-level counters of the last step (JSON lines)")
    ap.add_argument("--no-replan", action="store_true", help="keep the snake-order shares (no planning from the warm-up's batch times)")
    ap.add_argument("--contiguous", action="store_true", help="shares in the file's group order instead of the source-degree order")
    ap.add_argument("--opt", action="append", default=[], help="engine option key=value (experiments)")
    ap.add_argument("--alpha", type=int, default=0)
    ap.add_argument("--lanes", type=int, default=0)
    args = ap.parse_args()
    claim_stdout()
    if args.warmup < 3 and args.impl == "ours":
        log("note: fewer than 3 warm-up steps requested; the timing rules ask for >= 3")
    rank, local_rank, world = dist_env()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl, rank)
        return
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch with torchrun for --gpus > 1 (one process per GPU)")
    run_ours(args, wl, rank, local_rank, world)


if __name__ == "__main__":
    main()
