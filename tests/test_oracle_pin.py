"""Pins the CPU oracle (oracle/oracle.c) to the reference's own code: against the golden
vectors that oracle/_ref/main_host produced (tests/golden/make_golden.py) and, where the
host build of the reference is present, against live runs of it."""
import os
import struct

import numpy as np
import pytest

from conftest import GOLDEN_CASES, golden_case_inputs


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle_matches_reference_golden(pkg, ob, expected, name):
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    case = expected[name]
    if "edge_checksum" in case:  # the regenerated graph is the one the reference saw
        assert int(uv.astype(np.int64).sum()) == case["edge_checksum"]
    res = ob.run(n, uv, offs, src)
    assert res.sums.tolist() == case["sums"]
    assert res.min_k + 1 == case["k_printed"]  # main.cu:409 prints minK + 1
    assert res.min_f == case["min_f"]
    # the all-threads-inside-one-BFS mode (bench.py's reference arm at Kronecker-26) and the
    # parallel CSR build are held to the same vectors
    csr = ob.Csr(n, uv, parallel=True)
    for threads in (1, 3):
        wide = ob.run_csr(csr, offs, src, threads=threads, wide=True)
        assert wide.sums.tolist() == case["sums"]
        assert np.array_equal(wide.reached_entries, res.reached_entries)


def test_known_answers_by_hand(ob):
    """SURVEY.md App. B, checked by hand: path 0-1-2-3-4, self-loop at 2, duplicate edge
    (0,1), vertices 5 and 6 isolated."""
    uv = np.array([(0, 1), (1, 2), (2, 3), (3, 4), (2, 2), (0, 1)], dtype=np.int32)
    csr = ob.Csr(7, uv)
    assert csr.row.tolist() == [0, 2, 5, 9, 11, 12, 12, 12]
    # record order inside a row, both directions, the self-loop twice (main.cu:114-115)
    assert csr.col[csr.row[2]:csr.row[3]].tolist() == [1, 3, 2, 2]
    assert ob.distances(csr, np.array([0], dtype=np.int32)).tolist() == [0, 1, 2, 3, 4, -1, -1]
    assert ob.distances(csr, np.array([100, -3], dtype=np.int32)).tolist() == [-1] * 7
    offs = np.array([0, 1, 1, 3, 4], dtype=np.int64)
    src = np.array([2, 100, -3, 0], dtype=np.int32)
    res = ob.run_csr(csr, offs, src)
    assert res.sums.tolist() == [6, 0, 0, 10]
    assert (res.min_k, res.min_f) == (1, 0)  # empty group wins, tie -> lower index
    assert res.reached_entries.tolist() == [12, 0, 0, 12]


def test_argmin_rules(ob):
    assert ob.argmin(np.array([], dtype=np.int64)) == (-1, -1)  # K = 0 (main.cu:379-380)
    assert ob.argmin(np.array([10, 10, 7, 7], dtype=np.int64)) == (2, 7)
    assert ob.argmin(np.array([5], dtype=np.int64)) == (0, 5)
    assert ob.argmin(np.array([3, 0, 0], dtype=np.int64)) == (1, 0)


def test_oracle_file_readers(ob, tmp_path, golden_dir):
    n, uv = ob.load_graph_file(os.path.join(golden_dir, "kat_g7.graph.bin"))
    assert n == 7 and uv.tolist() == [[0, 1], [1, 2], [2, 3], [3, 4], [2, 2], [0, 1]]
    offs, src = ob.load_query_file(os.path.join(golden_dir, "kat2.query.bin"))
    assert offs.tolist() == [0, 1, 1, 3, 4] and src.tolist() == [2, 100, -3, 0]
    offs, src = ob.load_query_file(os.path.join(golden_dir, "kat4.query.bin"))
    assert offs.tolist() == [0] and src.size == 0
    # extended container: K = 300 groups cannot be expressed with a uint8 count
    p = tmp_path / "ext.bin"
    with open(p, "wb") as f:
        f.write(b"\x00MSBFSQ2" + struct.pack("<I", 300))
        for i in range(300):
            f.write(struct.pack("<I", i % 3))
            f.write(np.arange(i % 3, dtype="<i4").tobytes())
    offs, src = ob.load_query_file(str(p))
    assert offs.size == 301 and offs[-1] == sum(i % 3 for i in range(300))
    with pytest.raises(OSError):
        ob.load_graph_file(str(tmp_path / "missing.bin"))


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(__file__), "..", "oracle", "_ref", "main_host")),
                    reason="host build of the reference not present")
@pytest.mark.parametrize("seed", [0, 1])
def test_oracle_matches_live_reference(pkg, ob, tmp_path, seed):
    """Fresh random inputs through the reference's own code (host build), full file for
    the argmin and K = 1 files for a sample of per-group sums."""
    rng = np.random.default_rng(seed)
    scale = 10
    n = 1 << scale
    uv = pkg.gen_edges_host(pkg.GEN_RMAT, scale, 0, 8 << scale, 100 + seed)
    groups = [rng.integers(-3, n + 3, size=int(rng.integers(0, 6))).astype(np.int32) for _ in range(40)]
    offs = np.concatenate([[0], np.cumsum([len(g) for g in groups])]).astype(np.int64)
    src = np.concatenate(groups).astype(np.int32) if offs[-1] else np.zeros(0, np.int32)
    gpath, qpath = str(tmp_path / "g.bin"), str(tmp_path / "q.bin")
    pkg.write_graph_file(gpath, n, uv)
    pkg.write_query_file(qpath, offs, src)
    res = ob.run(n, uv, offs, src)
    k, f, _ = ob.run_reference_host(gpath, qpath)
    assert (k, f) == (res.min_k + 1, res.min_f)
    k2, f2, _ = ob.run_reference_host(gpath, qpath, ranks=3)
    assert (k2, f2) == (k, f)
    for i in rng.choice(40, size=6, replace=False):
        q1 = str(tmp_path / "q1.bin")
        pkg.write_query_file(q1, np.array([0, len(groups[i])], dtype=np.int64), groups[i])
        _, f1, _ = ob.run_reference_host(gpath, q1)
        assert f1 == res.sums[i]


def test_oracle_against_an_independent_bfs(ob):
    """A third opinion: scipy's unweighted shortest paths (multi-source = min over sources)."""
    import scipy.sparse as sp
    from scipy.sparse.csgraph import shortest_path

    rng = np.random.default_rng(7)
    for n, m in [(50, 60), (200, 150), (300, 1200)]:
        uv = rng.integers(0, n, size=(m, 2)).astype(np.int32)   # duplicates and self-loops included
        a = sp.coo_matrix((np.ones(m), (uv[:, 0], uv[:, 1])), shape=(n, n))
        dist = shortest_path(a + a.T, method="D", unweighted=True)
        csr = ob.Csr(n, uv)
        for _ in range(10):
            src = rng.integers(-2, n + 2, size=int(rng.integers(0, 5))).astype(np.int32)
            valid = src[(src >= 0) & (src < n)]
            want = np.full(n, -1, dtype=np.int64)
            if valid.size:
                d = dist[valid].min(axis=0)
                want = np.where(np.isfinite(d), d, -1).astype(np.int64)
            got = ob.distances(csr, src)
            assert got.tolist() == want.tolist()
            res = ob.run_csr(csr, np.array([0, src.size], dtype=np.int64), src)
            assert res.sums[0] == want[want >= 0].sum()
