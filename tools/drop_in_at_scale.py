"""Run the drop-in executable itself at the headline scale (SURVEY 8(d), config 3's launch line):

    mpirun -np G host/main -g graph.bin -q query.bin -gn G        G in {1, 2, 4, 8}

on a Kronecker scale-26 graph.bin (2^30 - 1 records, 8 GiB) and an extended-format query.bin
(4096 groups x 16 sources) written to /dev/shm, and keep the seven report lines, the
MSBFS_STATS line of every rank and the wall time per G.  Measurement helper, not product code.

    python tools/drop_in_at_scale.py --workload kron26 --ranks 1,2,4,8 --out gpurun_out/r2/dropin_kron26.json
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="kron26", choices=sorted(bench.WORKLOADS))
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--out", default="")
    ap.add_argument("--repeat", type=int, default=2, help="runs per rank count (the second one has a warm page cache)")
    args = ap.parse_args()
    import torch

    pkg = graft.load_package()
    wl = bench.WORKLOADS[args.workload]
    kind = bench.KINDS[wl["kind"]]
    n = pkg.gen_num_vertices(kind, wl["p0"], wl["p1"])
    g, q = os.path.join(args.dir, "graph.bin"), os.path.join(args.dir, "query.bin")
    t0 = time.time()
    uv = bench.make_edges_host(pkg, wl, kind)
    t1 = time.time()
    pkg.write_graph_file(g, n, uv)
    del uv
    offs, src = pkg.gen_queries_host(n, wl["groups"], wl["per_group"], bench.QUERY_SEED)
    pkg.write_query_file(q, offs, src)
    t2 = time.time()
    print(f"graph.bin {os.path.getsize(g) / 2**30:.2f} GiB generated in {t1 - t0:.1f} s, written in {t2 - t1:.1f} s; "
          f"query.bin {os.path.getsize(q)} bytes", flush=True)
    ngpu = torch.cuda.device_count()
    main_exe = os.path.join(pkg.PKG_DIR, "host", "main")
    mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
    results = {"workload": args.workload, "description": wl["desc"], "graph_bytes": os.path.getsize(g), "gpus_on_box": ngpu, "runs": []}
    env = dict(os.environ, MSBFS_STATS="1")
    for r in [int(x) for x in args.ranks.split(",")]:
        if r > ngpu:
            print(f"skipping np={r}: only {ngpu} GPUs", flush=True)
            continue
        for rep in range(args.repeat):
            t0 = time.time()
            res = subprocess.run([mpirun, "-np", str(r), main_exe, "-g", g, "-q", q, "-gn", str(r)], capture_output=True, text=True,
                                 env=env, timeout=1200)
            wall = time.time() - t0
            rec = {"np": r, "gn": r, "repeat": rep, "rc": res.returncode, "wall_s": round(wall, 3), "report": res.stdout.splitlines(),
                   "stats": [ln for ln in res.stderr.splitlines() if ln.startswith("[rank")],
                   "stderr_other": [ln for ln in res.stderr.splitlines() if not ln.startswith("[rank")][-5:]}
            for ln in res.stdout.splitlines():
                if ln.startswith("Query number"): rec["k_printed"] = int(ln.split(":")[1])
                elif ln.startswith("Minimum F value"): rec["min_f"] = int(ln.split(":")[1])
                elif ln.startswith("Preprocessing time"): rec["preprocessing_s"] = float(ln.split()[2])
                elif ln.startswith("Computation time"): rec["computation_s"] = float(ln.split()[2])
            if "computation_s" in rec:
                rec["queries_per_sec"] = wl["groups"] / rec["computation_s"]
            results["runs"].append(rec)
            print(json.dumps({k: v for k, v in rec.items() if k not in ("report", "stats")}), flush=True)
            for ln in rec["stats"]:
                print("   ", ln, flush=True)
    os.remove(g)
    os.remove(q)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
