"""Parity of the CUDA path with the reference, through the C ABI (ctypes -> libmsbfs.so):
golden vectors produced by the reference's own code, the CPU oracle on seeded inputs, and
size-independent properties at the larger BASELINE configs.  Bit-exact: everything on this
path is integer arithmetic (int64 sums)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_CASES, golden_case_inputs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine(pkg):
    eng = pkg.Engine(0)
    yield eng
    eng.close()


def _reset(eng):
    eng.set_option("reorder", 1)
    eng.set_option("lanes", 0)
    eng.set_option("direction", 0)
    eng.set_option("alpha", 40)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden_all_modes(pkg, ob, expected, engine, name):
    """Every golden case, every lane width, both forced directions and the heuristic."""
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    case = expected[name]
    engine.set_graph_edges(n, uv)
    try:
        for lanes in (0, 64, 128, 256, 512):
            for direction in (0, 1, 2):
                engine.set_option("lanes", lanes)
                engine.set_option("direction", direction)
                for reorder in ((1, 0) if lanes == 0 else (1,)):  # heaviest-first run order / caller's order
                    engine.set_option("reorder", reorder)
                    rep = engine.solve(offs, src)
                    assert rep.sums.tolist() == case["sums"], (lanes, direction, reorder)
                    assert rep.min_k + 1 == case["k_printed"] and rep.min_f == case["min_f"]
    finally:
        _reset(engine)


@pytest.mark.parametrize("name", ["kat1", "rmat12_ragged"])
def test_golden_through_file_loader(pkg, ob, expected, engine, tmp_path, name):
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    gpath = str(tmp_path / "g.bin")
    pkg.write_graph_file(gpath, n, uv)
    assert engine.load_graph_file(gpath) == (n, len(uv))
    o2, s2 = pkg.load_query_file(os.path.join(os.path.dirname(__file__), "golden", expected[name]["query_file"]))
    assert engine.run(o2, s2).tolist() == expected[name]["sums"]


def test_loader_errors(pkg, engine, tmp_path):
    with pytest.raises(pkg.MsbfsError) as e:
        engine.load_graph_file(str(tmp_path / "missing.bin"))
    assert e.value.code == 2 and "Could not open graph file" in e.value.message
    p = str(tmp_path / "bad.bin")
    pkg.write_graph_file(p, 4, np.array([(0, 1), (1, 7)], dtype=np.int32))  # id 7 >= n: UB in the reference
    with pytest.raises(pkg.MsbfsError) as e:
        engine.load_graph_file(p)
    assert e.value.code == 4
    raw = open(p, "rb").read()
    open(p, "wb").write(raw[:-4])
    with pytest.raises(pkg.MsbfsError) as e:
        engine.load_graph_file(p)
    assert e.value.code == 2
    with pytest.raises(pkg.MsbfsError):
        engine.set_option("lanes", 100)
    # a crafted 12-byte header whose record count wraps 12 + 8 m and 2 m: refused before any arithmetic
    import struct
    for m_bad in (1 << 62, (1 << 61) + 5, (1 << 39) + 1):
        open(p, "wb").write(struct.pack("<iq", 4, m_bad))
        with pytest.raises(pkg.MsbfsError) as e:
            engine.load_graph_file(p)
        assert e.value.code == 6, e.value.message
    # a refused file leaves the graph that was loaded before in place
    uv = np.array([(0, 1), (1, 2)], dtype=np.int32)
    engine.set_graph_edges(3, uv)
    pkg.write_graph_file(p, 3, np.array([(0, 9)], dtype=np.int32))
    with pytest.raises(pkg.MsbfsError):
        engine.load_graph_file(p)
    assert engine.run(np.array([0, 1], dtype=np.int64), np.array([0], dtype=np.int32)).tolist() == [3]


def test_empty_and_degenerate_graphs(pkg, ob, engine):
    offs = np.array([0, 1, 3], dtype=np.int64)
    src = np.array([0, 1, 5], dtype=np.int32)
    engine.set_graph_edges(6, np.zeros((0, 2), dtype=np.int32))  # no edges: every F is 0
    assert engine.run(offs, src).tolist() == [0, 0]
    engine.set_graph_edges(3, np.array([(1, 1)], dtype=np.int32))  # a lone self-loop
    assert engine.run(offs, src).tolist() == ob.run(3, np.array([(1, 1)], dtype=np.int32), offs, src).sums.tolist()
    assert engine.run(np.array([0], dtype=np.int64), np.zeros(0, np.int32)).size == 0  # K = 0
    assert pkg.argmin(np.zeros(0, np.int64)) == (-1, -1)


def _ragged_queries(rng, n, k, max_s):
    sizes = rng.integers(0, max_s + 1, size=k)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    src = rng.integers(-4, n + 4, size=int(offs[-1])).astype(np.int32)
    return offs, src


@pytest.mark.parametrize("kind,p0,p1,m,k,max_s", [
    ("rmat", 14, 0, 16 << 14, 300, 8),      # hubs with chunked rows, ragged groups, 2 batches
    ("kron", 13, 0, 16 << 13, 257, 3),      # one group past a full 256-lane batch
    ("er", 20000, 0, 40000, 65, 300),       # sparse ER (many components), S > 255
    ("lattice", 64, 50, None, 130, 2),      # high diameter
])
def test_random_inputs_vs_oracle(pkg, ob, engine, kind, p0, p1, m, k, max_s):
    kinds = {"rmat": pkg.GEN_RMAT, "kron": pkg.GEN_KRON, "er": pkg.GEN_ER, "lattice": pkg.GEN_LATTICE}
    if m is None:
        m = pkg.lattice_num_edges(p0, p1)
    n = pkg.gen_num_vertices(kinds[kind], p0, p1)
    uv = pkg.gen_edges_host(kinds[kind], p0, p1, m, 77)
    offs, src = _ragged_queries(np.random.default_rng(k), n, k, max_s)
    want = ob.run(n, uv, offs, src)
    engine.set_graph_edges(n, uv)
    try:
        for direction in (0, 1, 2):
            engine.set_option("direction", direction)
            got = engine.solve(offs, src)
            assert np.array_equal(got.sums, want.sums), direction
            assert (got.min_k, got.min_f) == (want.min_k, want.min_f)
        engine.set_option("direction", 0)
        engine.upload_queries(offs, src)
        assert np.array_equal(engine.reached_entries(), want.reached_entries)
    finally:
        _reset(engine)


def test_star_hub_and_offsets_not_from_zero(pkg, ob, engine):
    """One vertex with 6000 neighbours (6 row chunks), a tail path, and a query window
    that starts in the middle of the offsets array (how ranks pass their shard)."""
    hub = 0
    leaves = np.arange(1, 6001, dtype=np.int32)
    star = np.stack([np.full_like(leaves, hub), leaves], axis=1)
    path = np.stack([np.arange(6000, 6100, dtype=np.int32), np.arange(6001, 6101, dtype=np.int32)], axis=1)
    uv = np.concatenate([star, path]).astype(np.int32)
    n = 6200
    offs, src = _ragged_queries(np.random.default_rng(9), n, 40, 3)
    want = ob.run(n, uv, offs, src)
    engine.set_graph_edges(n, uv)
    for direction in (1, 2, 0):
        engine.set_option("direction", direction)
        assert np.array_equal(engine.run(offs, src), want.sums)
    _reset(engine)
    sums = np.empty(15, dtype=np.int64)
    import ctypes as C
    rc = pkg.lib().msbfs_run(engine._h, 15, C.c_void_p(offs[10:].ctypes.data), C.c_void_p(src.ctypes.data), C.c_void_p(sums.ctypes.data))
    assert rc == 0 and np.array_equal(sums, want.sums[10:25])


def test_device_generator_matches_host(pkg):
    import torch
    for kind, p0, p1, m in [(pkg.GEN_RMAT, 12, 0, 30000), (pkg.GEN_KRON, 12, 0, 30000),
                            (pkg.GEN_LATTICE, 33, 21, pkg.lattice_num_edges(33, 21)), (pkg.GEN_ER, 5000, 0, 30000)]:
        buf = torch.empty((m, 2), dtype=torch.int32, device="cuda:0")
        rc = pkg.lib().msbfs_gen_edges_device(0, kind, p0, p1, m, 5, buf.data_ptr())
        assert rc == 0
        assert np.array_equal(buf.cpu().numpy(), pkg.gen_edges_host(kind, p0, p1, m, 5))


def test_mid_size_rmat20_vs_oracle(pkg, ob, engine):
    """RMAT scale-20 built from device-resident records; 512 groups x 8 against the oracle."""
    import torch
    scale, k, s = 20, 512, 8
    n, m = 1 << scale, 16 << scale
    buf = torch.empty((m, 2), dtype=torch.int32, device="cuda:0")
    assert pkg.lib().msbfs_gen_edges_device(0, pkg.GEN_RMAT, scale, 0, m, 3, buf.data_ptr()) == 0
    engine.set_graph_edges_device(n, m, buf.data_ptr())
    uv = buf.cpu().numpy()
    del buf
    offs, src = pkg.gen_queries_host(n, k, s, 4)
    got = engine.solve(offs, src)
    want = ob.run(n, uv, offs, src)
    assert np.array_equal(got.sums, want.sums) and (got.min_k, got.min_f) == (want.min_k, want.min_f)
    st = engine.stats()
    assert st["levels_bottom_up"] > 0 and st["levels_top_down"] > 0 and st["kernel_launches"] > 0


def test_full_size_config2_properties(pkg, ob, engine):
    """BASELINE config 2 (RMAT scale-22, 1024 groups x 8) at full size: the oracle on a
    sample of groups, and properties that need no oracle — F does not depend on the lane
    width, on the traversal direction, or on the order / duplication of a group's sources."""
    import torch
    scale, k, s = 22, 1024, 8
    n, m = 1 << scale, 16 << scale
    buf = torch.empty((m, 2), dtype=torch.int32, device="cuda:0")
    assert pkg.lib().msbfs_gen_edges_device(0, pkg.GEN_RMAT, scale, 0, m, 1, buf.data_ptr()) == 0
    engine.set_graph_edges_device(n, m, buf.data_ptr())
    uv = buf.cpu().numpy()
    del buf
    offs, src = pkg.gen_queries_host(n, k, s, 2)
    base = engine.run(offs, src)
    try:
        engine.set_option("lanes", 64)
        assert np.array_equal(engine.run(offs[:129], src), base[:128])
        engine.set_option("lanes", 0)
        engine.set_option("direction", 2)
        assert np.array_equal(engine.run(offs[:257], src), base[:256])
    finally:
        _reset(engine)
    # permuted + duplicated sources give the same F
    g = src.reshape(k, s)[:256]
    twisted = np.concatenate([g[:, ::-1], g[:, :3]], axis=1).astype(np.int32)
    offs2 = (np.arange(257) * twisted.shape[1]).astype(np.int64)
    assert np.array_equal(engine.run(offs2, twisted.ravel()), base[:256])
    # oracle on a sample
    pick = np.array([0, 1, 255, 256, 511, 777, 1023])
    o3 = (np.arange(len(pick) + 1) * s).astype(np.int64)
    want = ob.run(n, uv, o3, src.reshape(k, s)[pick].ravel())
    assert np.array_equal(base[pick], want.sums)


def test_lane_major_engine_and_its_fallback(pkg, ob, engine):
    """Road-like graphs run on the second engine (csrc/lane_major.cu: a visited bitmap per group, a
    queue of (group, vertex) pairs).  Same sums as the oracle and as the bit-parallel engine, with and
    without the fixed-width row records, the traversed-edges pass included; and on a graph whose
    frontiers outgrow the queue the batch is handed back to the bit-parallel engine."""
    def slots(eng):
        return {p["name"]: p["launches"] for p in eng.kernel_profiles()}

    try:
        # ---- lattice (rows <= 4 entries) and a path (rows <= 2, 3000 levels), ragged groups with bad ids
        path_uv = np.stack([np.arange(2999), np.arange(1, 3000)], axis=1).astype(np.int32)
        for n, uv, k in ((64 * 50, pkg.gen_edges_host(pkg.GEN_LATTICE, 64, 50, pkg.lattice_num_edges(64, 50), 5), 130),
                         (3000, path_uv, 70)):
            offs, src = _ragged_queries(np.random.default_rng(n), n, k, 3)
            want = ob.run(n, uv, offs, src)
            engine.set_graph_edges(n, uv)
            for lane_major, ell in ((-1, 1), (1, 0), (0, 1)):
                engine.set_option("lane_major", lane_major)
                engine.set_option("lm_ell", ell)
                assert np.array_equal(engine.run(offs, src), want.sums), (n, lane_major, ell)
                assert (slots(engine)["lane_major"] > 0) == (lane_major != 0)
                engine.upload_queries(offs, src)
                assert np.array_equal(engine.reached_entries(), want.reached_entries), (n, lane_major, ell)
        # ---- sparse random graph, rows <= 16 entries but exploding frontiers: the queue (2 x the non-isolated
        #      vertices, at least 65,536 entries) overflows and the bit-parallel engine redoes the batch
        n, m = 60000, 150000
        uv = pkg.gen_edges_host(pkg.GEN_ER, n, 0, m, 9)
        offs, src = pkg.gen_queries_host(n, 96, 3, 4)
        want = ob.run(n, uv, offs, src)
        engine.set_option("lane_major", 1)
        engine.set_option("streak_max_deg", 256)
        engine.set_graph_edges(n, uv)
        assert np.array_equal(engine.run(offs, src), want.sums)
        got = slots(engine)
        assert got["lane_major"] >= 1 and got["count"] > 0, got   # tried, overflowed, redone by the other engine
    finally:
        _reset(engine)
        engine.set_option("lane_major", -1)
        engine.set_option("lm_ell", 1)


def test_level_log_and_kernel_profiles(pkg, ob, expected, engine):
    """The measurement side of the ABI: per-level records and per-kernel accounting add up."""
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, "rmat16_64x4")
    engine.set_graph_edges(n, uv)
    engine.set_option("profile", 1)
    try:
        sums = engine.run(offs, src)
    finally:
        engine.set_option("profile", 0)
    assert sums.tolist() == expected["rmat16_64x4"]["sums"]
    st, log, prof = engine.stats(), engine.level_log(), engine.kernel_profiles()
    assert log[0]["direction"] == 0 and log[0]["level"] == -1          # source injection record
    assert len(log) == st["levels"] + st["batches"]                    # one record per level + injection
    assert sum(r["scanned"] for r in log) == st["scanned_top_down"] + st["scanned_bottom_up"]
    assert sum(p["scanned"] for p in prof) == sum(r["scanned"] for r in log)
    assert sum(p["launches"] for p in prof) == st["kernel_launches"]
    assert any(p["ms"] > 0 for p in prof) and st["alg_bytes"] > 0 and st["run_ms"] > 0


def test_nccl_min_allreduce_two_gpus(pkg, tmp_path):
    """msbfs_allreduce_min_u64 over a real ncclComm_t (needs two GPUs; skipped otherwise)."""
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    exe = str(tmp_path / "nccl_min_check")
    src_file = os.path.join(os.path.dirname(__file__), "native", "nccl_min_check.cpp")
    subprocess.run(["g++", "-std=c++17", "-O1", src_file, "-I", os.path.join(pkg.REPO_ROOT, "include"), "-L", pkg.PKG_DIR,
                    "-lmsbfs", "-ldl", "-lpthread", f"-Wl,-rpath,{pkg.PKG_DIR}", "-o", exe], check=True)
    res = subprocess.run([exe, "2"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and ("OK:" in res.stdout or "SKIP:" in res.stdout), res.stdout + res.stderr  # NCCL prints a banner first
