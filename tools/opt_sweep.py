"""Experiment helper (not part of the product): build one workload's graph once, then time the
resident traversal under several run-time option sets in the same process, check that every
set gives the same sums, and print per-batch and per-level kernel times.

usage: python tools/opt_sweep.py WORKLOAD [--groups N] [--steps K] "k=v,k=v" "base" ...
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402
import bench  # noqa: E402

SLOTS = ["td_heavy", "td_g32", "td_g8", "td_g2", "bu_seg", "bu_hub", "-", "bu_thread", "td_fin", "bu_hfin", "count", "book"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("workload")
    ap.add_argument("specs", nargs="+")
    ap.add_argument("--groups", type=int, default=0, help="use only the first N groups of the run order's file")
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--levels", action="store_true", help="print per-level kernel times of the first and last batch")
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    import torch

    pkg = graft.load_package()
    wl = bench.WORKLOADS[args.workload]
    kind = bench.KINDS[wl["kind"]]
    n = pkg.gen_num_vertices(kind, wl["p0"], wl["p1"])
    m = wl["m"]
    buf = torch.empty((m, 2), dtype=torch.int32, device="cuda:0")
    assert pkg.lib().msbfs_gen_edges_device(0, kind, wl["p0"], wl["p1"], m, bench.GRAPH_SEED, buf.data_ptr()) == 0
    eng = pkg.Engine(0)
    eng.set_option("profile", 1)
    eng.set_graph_edges_device(n, m, buf.data_ptr())
    del buf
    torch.cuda.empty_cache()
    offs, src = pkg.gen_queries_host(n, wl["groups"], wl["per_group"], bench.QUERY_SEED)
    if args.groups:
        offs = offs[: args.groups + 1]
        src = src[: offs[-1]]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
    ref_sums = None
    results = []
    defaults = {}
    for spec in args.specs:
        kvs = [kv.split("=") for kv in spec.split(",") if kv and kv != "base"]
        try:
            for k, v in kvs:
                eng.set_option(k, int(v))
            eng.upload_queries(offs, src)
            eng.run_resident()  # warm-up
            times = []
            for _ in range(args.steps):
                flush.zero_()
                torch.cuda.synchronize()
                eng.run_resident()
                times.append(eng.stats()["run_ms"])
            sums = eng.download_sums()
            same = True if ref_sums is None else bool(np.array_equal(sums, ref_sums))
            if ref_sums is None:
                ref_sums = sums.copy()
            prof = {p["name"]: round(p["ms"], 2) for p in eng.kernel_profiles() if p["launches"]}
            bms = [round(x, 2) for x in eng.batch_ms()]
            ms = float(np.mean(times))
            print(f"{args.workload} [{spec}] ms/step={ms:.2f} same_sums={same} batches={bms} kernels={prof}", flush=True)
            rec = {"spec": spec, "ms": ms, "same": same, "batch_ms": bms, "kernels": prof}
            if args.levels:
                log = eng.launch_log()
                nb = len(bms)
                for b in sorted({0, nb - 1}):
                    per = {}
                    for r in log:
                        if r["batch"] == b:
                            per.setdefault(r["level"], {}).setdefault(SLOTS[r["slot"]], 0.0)
                            per[r["level"]][SLOTS[r["slot"]]] += r["ms"]
                    txt = "; ".join(f"L{lv}: " + " ".join(f"{k}={v:.2f}" for k, v in sorted(d.items()) if v >= 0.05)
                                    for lv, d in sorted(per.items()) if sum(d.values()) >= 0.1)
                    print(f"    batch {b}: {txt}", flush=True)
                    rec[f"levels_batch{b}"] = {str(k): v for k, v in per.items()}
            results.append(rec)
        except Exception as e:  # keep sweeping
            print(f"{args.workload} [{spec}] FAILED {e}", flush=True)
        finally:
            for k, _ in kvs:  # back to defaults for the next spec
                try:
                    eng.set_option(k, defaults.get(k, {"col_width": 0, "profile": 1, "lanes": 0, "direction": 0, "reorder": 1,
                                                       "alpha": 40}.get(k, 0)))
                except Exception:
                    pass
    if args.json:
        with open(args.json, "w") as f:
            json.dump(results, f)
    assert all(r["same"] for r in results), "option sets disagree on the sums"


if __name__ == "__main__":
    main()
