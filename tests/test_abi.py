"""CPU-side checks of the C-ABI library: it loads without a GPU, exports every symbol
include/msbfs.h declares, its host-only entry points behave like the reference's host
code, and the compute entry points fail loudly (no fallback) when there is no device."""
import os
import re
import struct

import numpy as np
import pytest


def declared_symbols(header_path):
    text = open(header_path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(msbfs_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    names = declared_symbols(pkg.HEADER_PATH)
    assert len(names) >= 25
    lib = pkg.lib()
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/msbfs.h but not exported"
    assert set(names) == set(pkg.SIGNATURES), "python binding table and header disagree"


def test_library_has_sm100a_code(pkg):
    import subprocess
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "--list-elf", pkg.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in out.stdout


def test_no_gpu_fails_loudly(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.MsbfsError) as e:
        pkg.Engine(0)
    assert "no CPU fallback" in str(e.value)


def test_argmin_matches_reference_rule(pkg, ob):
    rng = np.random.default_rng(3)
    for _ in range(200):
        k = int(rng.integers(0, 12))
        sums = rng.integers(0, 5, size=k).astype(np.int64)
        assert pkg.argmin(sums) == ob.argmin(sums)
    assert pkg.argmin(np.zeros(0, np.int64)) == (-1, -1)


def test_packed_argmin(pkg):
    sums = np.array([10, 10, 7, 7], dtype=np.int64)
    assert pkg.unpack_argmin(pkg.argmin_packed(sums, 0)) == (2, 7)
    assert pkg.unpack_argmin(pkg.argmin_packed(sums, 1000)) == (1002, 7)
    assert pkg.argmin_packed(np.zeros(0, np.int64), 5) == (1 << 64) - 1
    assert pkg.unpack_argmin((1 << 64) - 1) == (-1, -1)
    # ties across shards: the smaller global id must win under a plain unsigned min
    a = pkg.argmin_packed(np.array([5, 3], dtype=np.int64), 0)
    b = pkg.argmin_packed(np.array([3, 9], dtype=np.int64), 2)
    assert pkg.unpack_argmin(min(a, b)) == (1, 3)
    with pytest.raises(pkg.MsbfsError):
        pkg.argmin_packed(np.array([1 << 40], dtype=np.int64), 0)
    with pytest.raises(pkg.MsbfsError):
        pkg.argmin_packed(np.array([1], dtype=np.int64), (1 << 24) - 1 + 1)


def test_query_file_reference_format_bytes(pkg, ob, tmp_path, golden_dir):
    """The writer emits the reference layout byte for byte (main.cu:143-161) and the
    reader agrees with the oracle's reader on the committed files."""
    offs = np.array([0, 1, 1, 3, 4], dtype=np.int64)
    src = np.array([2, 100, -3, 0], dtype=np.int32)
    p = str(tmp_path / "q.bin")
    pkg.write_query_file(p, offs, src)
    assert open(p, "rb").read() == open(os.path.join(golden_dir, "kat2.query.bin"), "rb").read()
    for name in os.listdir(golden_dir):
        if name.endswith(".query.bin"):
            a = pkg.load_query_file(os.path.join(golden_dir, name))
            b = ob.load_query_file(os.path.join(golden_dir, name))
            assert a[0].tolist() == b[0].tolist() and a[1].tolist() == b[1].tolist()
    # K = 0 is the single byte 0x00
    pkg.write_query_file(p, np.array([0], dtype=np.int64), np.zeros(0, np.int32))
    assert open(p, "rb").read() == b"\x00"
    o, s = pkg.load_query_file(p)
    assert o.tolist() == [0] and s.size == 0


def test_query_file_extended_format(pkg, ob, tmp_path):
    rng = np.random.default_rng(0)
    sizes = rng.integers(0, 400, size=700)
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    src = rng.integers(-10, 1 << 20, size=int(offs[-1])).astype(np.int32)
    p = str(tmp_path / "ext.bin")
    pkg.write_query_file(p, offs, src)
    raw = open(p, "rb").read()
    assert raw[:8] == b"\x00MSBFSQ2" and struct.unpack("<I", raw[8:12])[0] == 700
    o, s = pkg.load_query_file(p)
    assert o.tolist() == offs.tolist() and s.tolist() == src.tolist()
    o2, s2 = ob.load_query_file(p)
    assert o2.tolist() == offs.tolist() and s2.tolist() == src.tolist()
    with pytest.raises(pkg.MsbfsError):
        pkg.load_query_file(str(tmp_path / "missing.bin"))
    open(p, "wb").write(raw[:-3])
    with pytest.raises(pkg.MsbfsError):
        pkg.load_query_file(p)


def test_graph_file_writer_layout(pkg, ob, tmp_path, golden_dir):
    uv = np.array([(0, 1), (1, 2), (2, 3), (3, 4), (2, 2), (0, 1)], dtype=np.int32)
    p = str(tmp_path / "g.bin")
    pkg.write_graph_file(p, 7, uv)
    assert open(p, "rb").read() == open(os.path.join(golden_dir, "kat_g7.graph.bin"), "rb").read()
    n, back = ob.load_graph_file(p)
    assert n == 7 and back.tolist() == uv.tolist()


def test_generators_are_deterministic_and_in_range(pkg):
    for kind, p0, p1, m in [(pkg.GEN_RMAT, 10, 0, 5000), (pkg.GEN_KRON, 10, 0, 5000),
                            (pkg.GEN_LATTICE, 9, 7, pkg.lattice_num_edges(9, 7)), (pkg.GEN_ER, 1000, 0, 5000)]:
        n = pkg.gen_num_vertices(kind, p0, p1)
        a = pkg.gen_edges_host(kind, p0, p1, m, 42)
        b = pkg.gen_edges_host(kind, p0, p1, m, 42)
        assert np.array_equal(a, b) and a.min() >= 0 and a.max() < n
        if kind != pkg.GEN_LATTICE:
            assert not np.array_equal(a, pkg.gen_edges_host(kind, p0, p1, m, 43))
    # lattice: every vertex has 2..4 neighbours, no duplicates
    uv = pkg.gen_edges_host(pkg.GEN_LATTICE, 9, 7, pkg.lattice_num_edges(9, 7), 0)
    deg = np.bincount(uv.ravel(), minlength=63)
    assert deg.min() == 2 and deg.max() == 4 and len({tuple(e) for e in uv.tolist()}) == len(uv)
    # kron is rmat under a bijection of the labels
    r = pkg.gen_edges_host(pkg.GEN_RMAT, 10, 0, 5000, 42)
    k = pkg.gen_edges_host(pkg.GEN_KRON, 10, 0, 5000, 42)
    mapping = {}
    for a_, b_ in zip(r.ravel().tolist(), k.ravel().tolist()):
        assert mapping.setdefault(a_, b_) == b_
    assert len(set(mapping.values())) == len(mapping)
    offs, src = pkg.gen_queries_host(1000, 7, 5, 1)
    assert offs.tolist() == [0, 5, 10, 15, 20, 25, 30, 35] and src.min() >= 0 and src.max() < 1000


def test_shard_ranges_cover_everything(pkg):
    for k in [0, 1, 255, 256, 257, 1024, 4096, 5000]:
        for world in [1, 2, 3, 4, 8]:
            covered = []
            for r in range(world):
                a, b = pkg.shard_range(k, r, world)
                assert 0 <= a <= b <= k
                covered += list(range(a, b))
            assert covered == list(range(k))


def test_share_planning_and_explicit_id_packing(pkg, ob):
    """plan_shares: every group lands on exactly one rank, batches keep the heaviest-first
    order; argmin_packed_ids + a plain min over ranks gives the reference's argmin."""
    rng = np.random.default_rng(11)
    for k, world in [(0, 2), (1, 3), (700, 2), (4096, 8), (4096, 4), (1000, 8), (5000, 3)]:
        costs = rng.integers(0, 50, size=k).astype(np.int64)
        share, lanes = pkg.plan_shares(costs, world)
        assert len(share) == world
        allg = np.concatenate(share) if k else np.zeros(0, np.int64)
        assert sorted(allg.tolist()) == list(range(k))
        if world > 1 and k:
            assert lanes in (64, 128, 256, 512)
            order = pkg.deal_groups(costs, 0, 1).tolist()
            pos = {g: i for i, g in enumerate(order)}
            for sh in share:   # inside a batch the global order is kept
                for b in range(0, len(sh), lanes):
                    p = [pos[g] for g in sh[b:b + lanes].tolist()]
                    assert p == list(range(p[0], p[0] + len(p)))
        sums = rng.integers(0, 6, size=k).astype(np.int64)
        packed = [pkg.argmin_packed_ids(sums[sh], sh) for sh in share]
        assert pkg.unpack_argmin(min(packed)) == ob.argmin(sums)
    # planning from measured batch cost: longest batch first onto the least-loaded rank
    costs = np.arange(4096, 0, -1).astype(np.int64)           # run order = file order
    measured = np.array([13, 14, 15, 16, 18, 20, 23, 26.6] * 2)  # 16 batches of 256
    share, lanes = pkg.plan_shares(costs, 8, lanes=256, batch_cost=measured)
    assert lanes == 256 and sorted(np.concatenate(share).tolist()) == list(range(4096))
    loads = [sum(measured[g // 256] for g in sh[::256].tolist()) for sh in share]
    snake, _ = pkg.plan_shares(costs, 8, lanes=256)
    loads_snake = [sum(measured[g // 256] for g in sh[::256].tolist()) for sh in snake]
    assert max(loads) <= max(loads_snake) and max(loads) < 1.15 * sum(measured) / 8
    # one batch per rank: 512 lanes when the batches weigh about the same (no hubs), 256 when they do not
    flat = 1000 + (np.arange(4096) % 7).astype(np.int64)
    assert pkg.plan_shares(flat, 8)[1] == 512 and pkg.plan_shares(flat, 2)[1] == 512
    skewed = (np.arange(4096, 0, -1).astype(np.int64)) ** 3
    assert pkg.plan_shares(skewed, 8)[1] == 256 and pkg.plan_shares(skewed, 4)[1] == 512  # two batches per rank at 4
    assert pkg.lib().msbfs_plan_lanes(4096, 1) == 0 and pkg.lib().msbfs_plan_lanes(4096, 8) == 256
    assert pkg.lib().msbfs_plan_lanes(4096, 2) == 512 and pkg.lib().msbfs_plan_lanes(100, 8) == 64
    offs = np.array([0, 2, 2, 5, 6], dtype=np.int64)
    src = np.arange(6, dtype=np.int32)
    o2, s2 = pkg.gather_groups(offs, src, np.array([2, 0, 1]))
    assert o2.tolist() == [0, 3, 5, 5] and s2.tolist() == [2, 3, 4, 0, 1]


def test_query_file_round_trip_fuzz(pkg, ob, tmp_path):
    """Any group structure survives write -> read, in whichever container the writer picks, and
    the engine's reader and the oracle's reader agree byte for byte."""
    from hypothesis import given, settings, strategies as st

    groups_st = st.lists(st.lists(st.integers(-(2 ** 31), 2 ** 31 - 1), max_size=300), max_size=300)
    path = str(tmp_path / "fuzz.bin")

    @settings(max_examples=60, deadline=None)
    @given(groups_st, st.booleans())
    def check(groups, force_ext):
        offs = np.concatenate([[0], np.cumsum([len(g) for g in groups])]).astype(np.int64)
        src = np.array([x for g in groups for x in g], dtype=np.int64).astype(np.int32)
        pkg.write_query_file(path, offs, src, force_extended=force_ext)
        raw = open(path, "rb").read()
        extended = raw[:8] == b"\x00MSBFSQ2"
        assert extended == (force_ext or len(groups) > 255 or any(len(g) > 255 for g in groups))
        for reader in (pkg.load_query_file, ob.load_query_file):
            o, s = reader(path)
            assert o.tolist() == offs.tolist() and s.tolist() == src.tolist()

    check()


def test_plan_shares_properties(pkg):
    """msbfs_plan_shares on arbitrary weights / world sizes / widths: the run order is the stable
    heaviest-first permutation, every group has exactly one owner in range, a batch (a run of
    `lanes` consecutive groups of the order) is never split between ranks, and with measured
    costs no rank is left idle while another holds two batches more than it needs to."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=150, deadline=None)
    @given(st.lists(st.integers(0, 1 << 40), min_size=0, max_size=1500), st.integers(1, 9), st.sampled_from([0, 64, 128, 256, 512]),
           st.booleans(), st.integers(0, 1 << 30))
    def check(weights, world, lanes, with_costs, seed):
        w = np.array(weights, dtype=np.int64)
        k = int(w.size)
        width = lanes or (pkg.lib().msbfs_plan_lanes(k, world) or 512)
        costs = None
        if with_costs and lanes and k:
            nb = (k + lanes - 1) // lanes
            costs = np.random.default_rng(seed).uniform(1.0, 30.0, size=nb)
        share, out_lanes = pkg.plan_shares(w, world, lanes=lanes, batch_cost=costs)
        assert len(share) == world
        allg = np.concatenate(share) if k else np.zeros(0, np.int64)
        assert sorted(allg.tolist()) == list(range(k))
        if world == 1 or k == 0:
            return
        assert out_lanes in (64, 128, 256, 512) and (lanes == 0 or out_lanes == lanes)
        order = np.lexsort((np.arange(k), -w))
        pos = np.empty(k, np.int64); pos[order] = np.arange(k)
        owner = np.empty(k, np.int64)
        for r, sh in enumerate(share):
            owner[sh] = r
            assert np.all(np.diff(pos[sh]) > 0)            # a rank runs its groups in the global run order
        for b in range(0, k, out_lanes):                    # a batch has one owner
            assert len(set(owner[order[b:b + out_lanes]].tolist())) == 1
        nb = (k + out_lanes - 1) // out_lanes
        per_rank = np.bincount(owner[order[::out_lanes]], minlength=world)
        if costs is None:
            assert per_rank.max() - per_rank.min() <= 1     # snake order deals the batches out evenly
        elif nb >= world:
            assert per_rank.min() >= 1                      # longest-first never leaves a rank empty

    check()


def test_bench_sample_spread_is_a_permutation():
    """bench.py's parity sample: group ids taken in turn from the first, a middle and the last 512-group
    batch of the run order, then the rest — every group exactly once, heaviest batch first."""
    import importlib
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    bench = importlib.import_module("bench")
    rng = np.random.default_rng(5)
    for k in (1, 17, 512, 513, 4096):
        costs = rng.integers(0, 1000, size=k).astype(np.int64)
        ids = bench.spread_over_run_order(costs)
        assert sorted(ids.tolist()) == list(range(k))
        order = np.lexsort((np.arange(k), -costs))
        assert ids[0] == order[0]
        if k > 1024:
            nb = (k + 511) // 512
            assert ids[1] == order[(nb // 2) * 512] and ids[2] == order[(nb - 1) * 512]
