"""Reported baseline (not a target): the reference's OWN GPU program — the unmodified
/root/reference/main.cu compiled for sm_100a by oracle/Makefile (oracle/_ref/main_gpu) —
run on the same box on a BASELINE workload, next to the new engine's drop-in `main` on
the same files.  The reference's query format caps a file at 255 groups x 255 sources and
its CSR at 2^31 entries, so it gets the first `--groups` groups (<= 255) of the workload.

    python tools/ref_gpu_baseline.py --workload rmat22 --groups 64 [--ranks 1]

Prints one JSON line with both programs' own "Preprocessing time" / "Computation time"."""
import argparse, json, os, subprocess, sys, tempfile, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft
import bench

def run(cmd, timeout):
    t0 = time.time()
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    wall = time.time() - t0
    if res.returncode != 0:
        return {"error": res.stderr[-400:], "wall_s": wall}
    out = {}
    for ln in res.stdout.splitlines():
        if ln.startswith("Query number"): out["k_printed"] = int(ln.split(":")[1])
        elif ln.startswith("Minimum F value"): out["min_f"] = int(ln.split(":")[1])
        elif ln.startswith("Preprocessing time"): out["preprocessing_s"] = float(ln.split()[2])
        elif ln.startswith("Computation time"): out["computation_s"] = float(ln.split()[2])
    out["wall_s"] = wall
    return out

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="rmat16", choices=sorted(bench.WORKLOADS))
    ap.add_argument("--groups", type=int, default=64)
    ap.add_argument("--ranks", type=int, default=1)
    ap.add_argument("--timeout", type=float, default=900)
    args = ap.parse_args()
    pkg = graft.load_package()
    wl = bench.WORKLOADS[args.workload]
    kind = bench.KINDS[wl["kind"]]
    n = pkg.gen_num_vertices(kind, wl["p0"], wl["p1"])
    uv = bench.make_edges_host(pkg, wl, kind)
    per_group = min(wl["per_group"], 255)
    groups = min(args.groups, wl["groups"], 255)
    offs, src = pkg.gen_queries_host(n, wl["groups"], wl["per_group"], bench.QUERY_SEED)
    # first `groups` groups, at most 255 sources each (reference file format)
    import numpy as np
    sel = [src[offs[i]:offs[i] + per_group] for i in range(groups)]
    offs2 = np.concatenate([[0], np.cumsum([len(s) for s in sel])]).astype(np.int64)
    src2 = np.concatenate(sel).astype(np.int32)
    tmpdir = "/dev/shm" if os.path.isdir("/dev/shm") else None
    with tempfile.TemporaryDirectory(dir=tmpdir) as tmp:
        g, q = os.path.join(tmp, "graph.bin"), os.path.join(tmp, "query.bin")
        pkg.write_graph_file(g, n, uv)
        pkg.write_query_file(q, offs2, src2)
        mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
        ref = os.path.join(ROOT, "oracle", "_ref", "main_gpu")
        ours = os.path.join(pkg.PKG_DIR, "host", "main")
        tail = ["-g", g, "-q", q, "-gn", "1"]
        r_ours = run([mpirun, "-np", str(args.ranks), ours] + tail, args.timeout)
        r_ref = run([mpirun, "-np", str(args.ranks), ref] + tail, args.timeout) if os.path.exists(ref) else {"error": "oracle/_ref/main_gpu not built"}
    line = {"workload": args.workload, "description": wl["desc"], "groups": groups, "sources_per_group": per_group,
            "ranks": args.ranks, "reference_main_cu_sm100a": r_ref, "new_engine_main": r_ours}
    if "computation_s" in r_ref and "computation_s" in r_ours:
        line["queries_per_sec"] = {"reference": groups / r_ref["computation_s"], "new": groups / r_ours["computation_s"]}
        line["same_answer"] = (r_ref.get("k_printed"), r_ref.get("min_f")) == (r_ours.get("k_printed"), r_ours.get("min_f"))
    print(json.dumps(line), flush=True)

if __name__ == "__main__":
    main()
