"""Generates tests/golden/* by running the REFERENCE'S OWN SOURCE (oracle/_ref/main_host,
the unmodified /root/reference/main.cu compiled for the host by oracle/Makefile) on small
inputs.  Run in the build container, where /root/reference exists:

    python tests/golden/make_golden.py

The reference prints only the argmin (1-based k) and min F (main.cu:409-410), so each
per-group F is obtained by feeding it a query file that holds that single group (K = 1:
its "Minimum F value" line is then that group's F).  Committed outputs:
  kat*.graph.bin / *.query.bin   tiny explicit inputs
  <case>.query.bin               query files of the generated cases
  expected.json                  per-case parameters, per-group F, argmin, ref stdout lines
Graphs of the generated cases are NOT stored: they are re-created bit-identically from
(kind, p0, p1, m, seed) by the engine's host generator (csrc/gen_common.h).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as graft  # noqa: E402
import oracle_binding as ob  # noqa: E402

pkg = graft.load_package()


def write_query_ref_format(path, groups):
    """Plain reference format (main.cu:143-161), written independently of the engine."""
    with open(path, "wb") as f:
        f.write(bytes([len(groups)]))
        for g in groups:
            f.write(bytes([len(g)]))
            f.write(np.asarray(g, dtype="<i4").tobytes())


def write_graph_ref_format(path, n, uv):
    with open(path, "wb") as f:
        f.write(np.int32(n).tobytes())
        f.write(np.int64(len(uv)).tobytes())
        f.write(np.asarray(uv, dtype="<i4").tobytes())


def ref_case(graph_path, groups, tmp, ranks=(1,)):
    """Run the reference on the full query file and on each group alone."""
    qpath = os.path.join(tmp, "q_full.bin")
    write_query_ref_format(qpath, groups)
    k, f, out = ob.run_reference_host(graph_path, qpath)
    multi = {}
    for r in ranks:
        if r == 1:
            continue
        kr, fr, _ = ob.run_reference_host(graph_path, qpath, ranks=r)
        multi[str(r)] = [kr, fr]
    sums = []
    for i, g in enumerate(groups):
        q1 = os.path.join(tmp, "q_one.bin")
        write_query_ref_format(q1, [g])
        k1, f1, _ = ob.run_reference_host(graph_path, q1)
        assert k1 == 1
        sums.append(f1)
    lines = [ln for ln in out.splitlines()[2:5]]
    return {"k_printed": k, "min_f": f, "sums": sums, "report_lines_3_to_5": lines, "multi_rank": multi}


def main():
    if not os.path.exists(ob.REF_HOST):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)
    expected = {"_generated_by": "tests/golden/make_golden.py with oracle/_ref/main_host (reference source, host build)"}
    with tempfile.TemporaryDirectory() as tmp:
        # ---- known-answer tests on the 7-vertex graph of SURVEY.md App. B
        g7 = [(0, 1), (1, 2), (2, 3), (3, 4), (2, 2), (0, 1)]
        g7_path = os.path.join(HERE, "kat_g7.graph.bin")
        write_graph_ref_format(g7_path, 7, g7)
        kats = {
            "kat1": [[0], [2], [4], [5], [0, 4], [2, 2, 2]],
            "kat2": [[2], [], [100, -3], [0]],
            "kat3": [[0], [4], [1], [3]],
        }
        for name, groups in kats.items():
            write_query_ref_format(os.path.join(HERE, f"{name}.query.bin"), groups)
            res = ref_case(g7_path, groups, tmp, ranks=(1, 2, 3))
            res.update({"graph_file": "kat_g7.graph.bin", "query_file": f"{name}.query.bin", "groups": groups})
            expected[name] = res
        # K = 0 (main.cu:379-380, 409-410 -> prints 0 and -1)
        g3_path = os.path.join(HERE, "kat_g3.graph.bin")
        write_graph_ref_format(g3_path, 3, [(0, 1), (1, 2)])
        write_query_ref_format(os.path.join(HERE, "kat4.query.bin"), [])
        k, f, out = ob.run_reference_host(g3_path, os.path.join(HERE, "kat4.query.bin"))
        expected["kat4"] = {"graph_file": "kat_g3.graph.bin", "query_file": "kat4.query.bin", "groups": [],
                            "k_printed": k, "min_f": f, "sums": [], "report_lines_3_to_5": out.splitlines()[2:5],
                            "multi_rank": {}}

        # ---- generated cases
        rng = np.random.default_rng(20251106)
        gen_cases = {
            # config 1 of BASELINE.json: RMAT-16, 64 groups x 4 uniform sources
            "rmat16_64x4": dict(kind=pkg.GEN_RMAT, p0=16, p1=0, m=1 << 20, seed=1, groups="uniform", k=64, s=4, qseed=2,
                                ranks=(1, 2, 3)),
            # ragged groups, out-of-range ids, empty groups (SURVEY.md App. B second bullet)
            "rmat12_ragged": dict(kind=pkg.GEN_RMAT, p0=12, p1=0, m=1 << 16, seed=3, groups="ragged", k=48, ranks=(1, 2)),
            "kron12_200x3": dict(kind=pkg.GEN_KRON, p0=12, p1=0, m=1 << 16, seed=5, groups="uniform", k=200, s=3, qseed=6,
                                 ranks=(1,)),
            "lattice_48x40": dict(kind=pkg.GEN_LATTICE, p0=48, p1=40, m=pkg.lattice_num_edges(48, 40), seed=0,
                                  groups="uniform", k=70, s=2, qseed=8, ranks=(1,)),
            "er_4096": dict(kind=pkg.GEN_ER, p0=4096, p1=0, m=1 << 15, seed=9, groups="uniform", k=20, s=40, qseed=10,
                            ranks=(1,)),
        }
        for name, c in gen_cases.items():
            n = pkg.gen_num_vertices(c["kind"], c["p0"], c["p1"])
            uv = pkg.gen_edges_host(c["kind"], c["p0"], c["p1"], c["m"], c["seed"])
            gpath = os.path.join(tmp, f"{name}.graph.bin")
            write_graph_ref_format(gpath, n, uv)
            if c["groups"] == "uniform":
                offs, src = pkg.gen_queries_host(n, c["k"], c["s"], c["qseed"])
                groups = [src[offs[i]:offs[i + 1]].tolist() for i in range(c["k"])]
            else:
                groups = [rng.integers(-5, n + 5, size=int(rng.integers(0, 9))).tolist() for _ in range(c["k"])]
                groups[7] = []
                groups[11] = [-1, n, n + 3]
            write_query_ref_format(os.path.join(HERE, f"{name}.query.bin"), groups)
            res = ref_case(gpath, groups, tmp, ranks=c["ranks"])
            res.update({"gen": {k: int(v) for k, v in c.items() if k in ("kind", "p0", "p1", "m", "seed")}, "n": n,
                        "query_file": f"{name}.query.bin", "edge_checksum": int(uv.astype(np.int64).sum())})
            expected[name] = res
            print(name, "k =", res["k_printed"], "F =", res["min_f"], flush=True)
    with open(os.path.join(HERE, "expected.json"), "w") as f:
        json.dump(expected, f, indent=1)
    print("wrote expected.json")


if __name__ == "__main__":
    main()
