"""Differential fuzz through the C ABI: many small random graphs of different shapes, ragged query
sets with out-of-range and repeated sources, and a random engine configuration per case (lane width,
direction, run order, count mode, engine choice, row format, col load width) — every per-group sum
and the argmin must equal the CPU oracle's (oracle/oracle.c, pinned to the reference's code).
Integer work: bit-exact.  Seeds are fixed, so a failure names a reproducible case."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

OPTIONS = {
    "lanes": [0, 64, 128, 256, 512],
    "direction": [0, 0, 1, 2],
    "reorder": [1, 0],
    "count_missing": [1, 0],
    "lane_major": [-1, 0, 1],
    "lm_ell": [1, 0],
    "col_width": [0, 4, 8],
    "alpha": [40, 4, 400],
    "tail": [1, 0, 2, 2],
}
DEFAULTS = {"lanes": 0, "direction": 0, "reorder": 1, "count_missing": 1, "lane_major": -1, "lm_ell": 1, "col_width": 0, "alpha": 40, "tail": 1}


def _graph(pkg, rng, shape):
    if shape == "path":
        n = int(rng.integers(2, 4000))
        uv = np.stack([np.arange(n - 1), np.arange(1, n)], axis=1)
    elif shape == "star":
        n = int(rng.integers(3, 9000))
        uv = np.stack([np.zeros(n - 1, np.int64), np.arange(1, n)], axis=1)
        uv = np.concatenate([uv, rng.integers(0, n, size=(int(rng.integers(0, 50)), 2))])
    elif shape == "lattice":
        w, h = int(rng.integers(2, 70)), int(rng.integers(2, 70))
        n = w * h
        uv = pkg.gen_edges_host(pkg.GEN_LATTICE, w, h, pkg.lattice_num_edges(w, h), 1)
    elif shape == "sparse":   # many components, short rows: the lane-major engine and its overflow fallback
        n = int(rng.integers(10, 30000))
        uv = rng.integers(0, n, size=(int(n * rng.uniform(0.3, 2.5)), 2))
    elif shape == "dense":
        n = int(rng.integers(5, 600))
        uv = rng.integers(0, n, size=(int(n * rng.uniform(4, 40)), 2))
    else:  # rmat: hubs, chunked rows, isolated vertices
        scale = int(rng.integers(6, 14))
        n = 1 << scale
        uv = pkg.gen_edges_host(pkg.GEN_RMAT, scale, 0, int((1 << scale) * rng.integers(2, 24)), int(rng.integers(1, 1 << 30)))
    return n, np.ascontiguousarray(uv, dtype=np.int32)


def _queries(rng, n):
    k = int(rng.choice([0, 1, 3, 63, 64, 65, 130, 257, 600]))
    sizes = rng.integers(0, int(rng.choice([2, 5, 40])) + 1, size=k)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    src = rng.integers(-3, n + 3, size=int(offs[-1])).astype(np.int32)
    if src.size > 4:
        src[rng.integers(0, src.size, size=src.size // 4)] = src[0]  # repeated sources
    return offs, src


@pytest.mark.parametrize("shape", ["path", "star", "lattice", "sparse", "dense", "rmat"])
def test_random_cases_match_the_oracle(pkg, ob, shape):
    rng = np.random.default_rng(sum(map(ord, shape)))
    with pkg.Engine(0) as eng:
        for case in range(40):
            n, uv = _graph(pkg, rng, shape)
            offs, src = _queries(rng, n)
            want = ob.run(n, uv, offs, src)
            eng.set_graph_edges(n, uv)
            for trial in range(4):
                cfg = {k: (v[int(rng.integers(0, len(v)))] if trial else DEFAULTS[k]) for k, v in OPTIONS.items()}
                for k, v in cfg.items():
                    eng.set_option(k, v)
                got = eng.solve(offs, src)
                assert np.array_equal(got.sums, want.sums), (shape, case, n, uv.shape[0], offs.size - 1, cfg)
                assert (got.min_k, got.min_f) == (want.min_k, want.min_f), (shape, case, cfg)
                if trial == 1 and offs.size > 1:
                    eng.upload_queries(offs, src)
                    assert np.array_equal(eng.reached_entries(), want.reached_entries), (shape, case, cfg)
