"""Python face of the B200 multi-source-BFS engine: a thin ctypes binding over the
C ABI in ``include/msbfs.h`` (``libmsbfs.so``) plus ``Engine``, a small mirror of the
reference program's flow (load graph -> load queries -> F per group -> argmin,
/root/reference/main.cu:230-397) used by the tests and by ``bench.py``.

The directory name carries hyphens (it is the reference's repository name), so import
it with ``importlib.import_module`` — see ``__graft_entry__.load_package``.

There is no CPU fallback anywhere in this package: every compute call goes through the
CUDA library and raises ``MsbfsError`` when the library or a GPU is missing.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
LIB_PATH = os.environ.get("MSBFS_LIB", os.path.join(PKG_DIR, "libmsbfs.so"))  # override: A/B builds of the kernels
HEADER_PATH = os.path.join(REPO_ROOT, "include", "msbfs.h")

GEN_RMAT, GEN_KRON, GEN_LATTICE, GEN_ER = 0, 1, 2, 3


class MsbfsError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"msbfs error {code}: {message}")
        self.code = code
        self.message = message


class Stats(C.Structure):
    """struct msbfs_stats (include/msbfs.h)."""

    _fields_ = [
        ("n", C.c_int32), ("n_active", C.c_int32), ("m", C.c_int64), ("csr_entries", C.c_int64),
        ("lanes", C.c_int32), ("batches", C.c_int32), ("levels", C.c_int64),
        ("levels_top_down", C.c_int64), ("levels_bottom_up", C.c_int64),
        ("scanned_top_down", C.c_int64), ("scanned_bottom_up", C.c_int64),
        ("frontier_vertices", C.c_int64), ("candidate_vertices", C.c_int64),
        ("alg_bytes", C.c_int64), ("kernel_launches", C.c_int64),
        ("build_ms", C.c_double), ("run_ms", C.c_double), ("bottom_up_ms", C.c_double),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


class KernelProfile(C.Structure):
    """struct msbfs_kernel_profile (include/msbfs.h)."""

    _fields_ = [
        ("name", C.c_char * 32), ("launches", C.c_int64), ("ms", C.c_double), ("alg_bytes", C.c_int64),
        ("scanned", C.c_int64), ("words_fetched", C.c_int64), ("vertices_examined", C.c_int64),
        ("words_written", C.c_int64),
    ]

    def as_dict(self) -> dict:
        d = {name: getattr(self, name) for name, _ in self._fields_}
        d["name"] = self.name.decode()
        return d


class LaunchRecord(C.Structure):
    """struct msbfs_launch_record (include/msbfs.h)."""

    _fields_ = [("batch", C.c_int32), ("level", C.c_int32), ("slot", C.c_int32), ("ms", C.c_float)]


class LevelRecord(C.Structure):
    """struct msbfs_level_record (include/msbfs.h)."""

    _fields_ = [
        ("level", C.c_int32), ("direction", C.c_int32), ("frontier_vertices", C.c_uint64),
        ("frontier_entries", C.c_uint64), ("unexplored_entries", C.c_int64), ("scanned", C.c_uint64),
        ("words_fetched", C.c_uint64), ("vertices_examined", C.c_uint64), ("words_written", C.c_uint64),
    ]


_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")

# name -> (restype, argtypes).  Every symbol declared in include/msbfs.h is listed here;
# tests/test_abi.py checks the header and this table against the built library.
SIGNATURES = {
    "msbfs_last_error": (C.c_char_p, []),
    "msbfs_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "msbfs_destroy": (C.c_int, [C.c_void_p]),
    "msbfs_set_option": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64]),
    "msbfs_load_graph_file": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "msbfs_set_graph_edges": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "msbfs_set_graph_edges_device": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p]),
    "msbfs_export_graph": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "msbfs_import_graph": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "msbfs_run": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "msbfs_upload_queries": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "msbfs_run_resident": (C.c_int, [C.c_void_p]),
    "msbfs_download_sums": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "msbfs_reached_entries": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32]),
    "msbfs_vertex_degrees": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "msbfs_group_weights": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "msbfs_small_run_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "msbfs_argmin": (C.c_int, [C.c_int32, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "msbfs_argmin_packed": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_uint64)]),
    "msbfs_argmin_packed_ids": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_uint64)]),
    "msbfs_unpack_argmin": (C.c_int, [C.c_uint64, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]),
    "msbfs_allreduce_min_u64": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]),
    "msbfs_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "msbfs_nccl_comm_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "msbfs_nccl_comm_destroy": (C.c_int, [C.c_void_p]),
    "msbfs_load_graph_file_sharded": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32),
                                                C.POINTER(C.c_int64)]),
    "msbfs_plan_lanes": (C.c_int, [C.c_int32, C.c_int32]),
    "msbfs_plan_shares": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_int32)]),
    "msbfs_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "msbfs_get_level_log": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]),
    "msbfs_get_launch_log": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]),
    "msbfs_get_batch_ms": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_int32)]),
    "msbfs_num_kernel_slots": (C.c_int, []),
    "msbfs_get_kernel_profile": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(KernelProfile)]),
    "msbfs_load_query_file": (C.c_int, [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "msbfs_write_query_file": (C.c_int, [C.c_char_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_int]),
    "msbfs_write_graph_file": (C.c_int, [C.c_char_p, C.c_int32, C.c_int64, C.c_void_p]),
    "msbfs_free": (None, [C.c_void_p]),
    "msbfs_gen_num_vertices": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]),
    "msbfs_gen_edges_host": (C.c_int, [C.c_int, C.c_int32, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p]),
    "msbfs_gen_edges_device": (C.c_int, [C.c_int, C.c_int, C.c_int32, C.c_int32, C.c_int64, C.c_uint64, C.c_void_p]),
    "msbfs_gen_queries_host": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_uint64, C.c_void_p, C.c_void_p]),
}

_lib = None


def build(verbose: bool = False) -> None:
    """Compile libmsbfs.so, host/main and host/mpirun in-tree (nvcc, sm_100a)."""
    res = subprocess.run(["make", "-C", PKG_DIR, "all"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise RuntimeError("building libmsbfs.so failed")


def lib() -> C.CDLL:
    """The loaded C-ABI library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MsbfsError(-1, f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise MsbfsError(rc, lib().msbfs_last_error().decode("utf-8", "replace"))


def _ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


# --------------------------------------------------------------------- host helpers

def gen_num_vertices(kind: int, p0: int, p1: int = 0) -> int:
    n = C.c_int32()
    _check(lib().msbfs_gen_num_vertices(kind, p0, p1, C.byref(n)))
    return n.value


def lattice_num_edges(w: int, h: int) -> int:
    return h * (w - 1) + (h - 1) * w


def gen_edges_host(kind: int, p0: int, p1: int, m: int, seed: int) -> np.ndarray:
    """(m, 2) int32 edge records, byte-identical to the device generator."""
    uv = np.empty((m, 2), dtype=np.int32)
    _check(lib().msbfs_gen_edges_host(kind, p0, p1, m, seed, _ptr(uv)))
    return uv


def gen_queries_host(n: int, groups: int, per_group: int, seed: int):
    offs = np.empty(groups + 1, dtype=np.int64)
    src = np.empty(max(groups * per_group, 1), dtype=np.int32)
    _check(lib().msbfs_gen_queries_host(n, groups, per_group, seed, _ptr(offs), _ptr(src)))
    return offs, src[: groups * per_group]


def write_graph_file(path: str, n: int, uv: np.ndarray) -> None:
    uv = np.ascontiguousarray(uv, dtype=np.int32).reshape(-1, 2)
    _check(lib().msbfs_write_graph_file(path.encode(), n, uv.shape[0], _ptr(uv)))


def write_query_file(path: str, offsets: np.ndarray, sources: np.ndarray, force_extended: bool = False) -> None:
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    sources = np.ascontiguousarray(sources, dtype=np.int32)
    if sources.size == 0:
        sources = np.zeros(1, dtype=np.int32)
    _check(lib().msbfs_write_query_file(path.encode(), offsets.size - 1, _ptr(offsets), _ptr(sources), int(force_extended)))


def load_query_file(path: str):
    k = C.c_int32()
    po, ps = C.c_void_p(), C.c_void_p()
    _check(lib().msbfs_load_query_file(path.encode(), C.byref(k), C.byref(po), C.byref(ps)))
    try:
        offs = np.ctypeslib.as_array(C.cast(po, C.POINTER(C.c_int64)), shape=(k.value + 1,)).copy()
        total = int(offs[-1])
        src = (np.ctypeslib.as_array(C.cast(ps, C.POINTER(C.c_int32)), shape=(max(total, 1),)).copy())[:total]
    finally:
        lib().msbfs_free(po)
        lib().msbfs_free(ps)
    return offs, src


def argmin(sums: np.ndarray):
    """(min_k, min_f) with the reference's rule (main.cu:379-397); (-1, -1) for K = 0."""
    sums = np.ascontiguousarray(sums, dtype=np.int64)
    k, f = C.c_int32(), C.c_int64()
    _check(lib().msbfs_argmin(sums.size, _ptr(sums) if sums.size else None, C.byref(k), C.byref(f)))
    return k.value, f.value


def argmin_packed(sums: np.ndarray, first_qid: int) -> int:
    sums = np.ascontiguousarray(sums, dtype=np.int64)
    out = C.c_uint64()
    _check(lib().msbfs_argmin_packed(_ptr(sums) if sums.size else None, first_qid, sums.size, C.byref(out)))
    return out.value


def argmin_packed_ids(sums: np.ndarray, query_ids: np.ndarray) -> int:
    sums = np.ascontiguousarray(sums, dtype=np.int64)
    ids = np.ascontiguousarray(query_ids, dtype=np.int32)
    out = C.c_uint64()
    _check(lib().msbfs_argmin_packed_ids(_ptr(sums) if sums.size else None, _ptr(ids) if ids.size else None, sums.size, C.byref(out)))
    return out.value


def unpack_argmin(packed: int):
    k, f = C.c_int32(), C.c_int64()
    _check(lib().msbfs_unpack_argmin(C.c_uint64(packed), C.byref(k), C.byref(f)))
    return k.value, f.value


def shard_range(num_groups: int, rank: int, world: int, lanes: int = 512):
    """Contiguous block of groups owned by `rank` (replaces the round-robin
    `kidx = rank; kidx += world_size` of main.cu:305).  Any partition gives the same argmin.
    Whole `lanes`-wide batches per rank keep every GPU's lane words full; when there are
    fewer batches than ranks the groups are split evenly in multiples of 64 lanes instead,
    so that no GPU idles."""
    batches = (num_groups + lanes - 1) // lanes
    if batches >= world:
        b0 = (batches * rank) // world
        b1 = (batches * (rank + 1)) // world
        return min(b0 * lanes, num_groups), min(b1 * lanes, num_groups)
    per = ((num_groups + world - 1) // world + 63) // 64 * 64
    lo = min(rank * per, num_groups)
    return lo, min(lo + per, num_groups)


def deal_groups(costs: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Group indices sorted heaviest first (ties by index), every world-th starting at `rank`;
    ``deal_groups(costs, 0, 1)`` is the whole order.  With costs = sum of a group's sources'
    degrees the order ranks groups by how fast their BFS reaches the hubs; batches cut from
    consecutive runs of it hold lanes that advance together, which is what lets the bottom-up
    kernels stop early for the whole batch.  Any assignment gives the same sums and argmin."""
    order = np.lexsort((np.arange(costs.size), -np.asarray(costs, dtype=np.int64)))
    return order[rank::world]


def plan_shares(costs: np.ndarray, world: int, lanes: int = 0, batch_cost=None):
    """(share, lanes) from msbfs_plan_shares — the one planner, shared with host/main.cpp:
    share[r] = group indices rank r runs, in run order; `lanes` = batch width to ask the
    engine for (0 = its own choice).  `batch_cost`: measured cost per batch of the same plan
    (batch order), if there is one; without it batches go out in snake order."""
    costs = np.ascontiguousarray(costs, dtype=np.int64)
    k = int(costs.size)
    order = np.zeros(max(k, 1), dtype=np.int32)
    owner = np.zeros(max(k, 1), dtype=np.int32)
    lanes_out = C.c_int32()
    bc = None if batch_cost is None else np.ascontiguousarray(batch_cost, dtype=np.float64)
    _check(lib().msbfs_plan_shares(k, _ptr(costs) if k else None, world, lanes, _ptr(bc) if bc is not None and bc.size else None,
                                   0 if bc is None else int(bc.size), _ptr(order), _ptr(owner), C.byref(lanes_out)))
    order, owner = order[:k].astype(np.int64), owner[:k]
    share = [order[owner[order] == r] for r in range(world)]
    return share, lanes_out.value


def gather_groups(offsets: np.ndarray, sources: np.ndarray, index: np.ndarray):
    """(offsets, sources) of the groups `index`, in that order."""
    sizes = (offsets[1:] - offsets[:-1])[index]
    new_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    parts = [sources[offsets[i]:offsets[i + 1]] for i in index]
    new_src = np.concatenate(parts).astype(np.int32) if len(parts) and new_off[-1] else np.zeros(0, np.int32)
    return new_off, new_src


# --------------------------------------------------------------------------- engine

@dataclass
class Report:
    """What the reference prints (main.cu:403-414), as values."""
    min_k: int      # 0-based; the report line shows min_k + 1
    min_f: int
    sums: np.ndarray


class Engine:
    """One handle on one GPU.  Mirrors main(): device setup -> graph -> queries -> F."""

    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        _check(lib().msbfs_create(device, C.byref(self._h)))
        self.device = device
        self.n = 0
        self.m = 0

    def close(self) -> None:
        if self._h:
            lib().msbfs_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key: str, value: int) -> None:
        _check(lib().msbfs_set_option(self._h, key.encode(), value))

    def load_graph_file(self, path: str):
        n, m = C.c_int32(), C.c_int64()
        _check(lib().msbfs_load_graph_file(self._h, path.encode(), C.byref(n), C.byref(m)))
        self.n, self.m = n.value, m.value
        return self.n, self.m

    def set_graph_edges(self, n: int, uv: np.ndarray) -> None:
        uv = np.ascontiguousarray(uv, dtype=np.int32).reshape(-1, 2)
        _check(lib().msbfs_set_graph_edges(self._h, n, uv.shape[0], _ptr(uv) if uv.size else None))
        self.n, self.m = n, uv.shape[0]

    def set_graph_edges_device(self, n: int, m: int, device_ptr: int) -> None:
        _check(lib().msbfs_set_graph_edges_device(self._h, n, m, C.c_void_p(device_ptr)))
        self.n, self.m = n, m

    def run(self, offsets: np.ndarray, sources: np.ndarray) -> np.ndarray:
        """F of every group; host buffers in, host buffer out (H2D and D2H inside)."""
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        sources = np.ascontiguousarray(sources, dtype=np.int32)
        k = offsets.size - 1
        sums = np.empty(max(k, 1), dtype=np.int64)
        src_ptr = _ptr(sources) if sources.size else None
        _check(lib().msbfs_run(self._h, k, _ptr(offsets), src_ptr, _ptr(sums)))
        return sums[:k]

    def upload_queries(self, offsets: np.ndarray, sources: np.ndarray) -> None:
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        sources = np.ascontiguousarray(sources, dtype=np.int32)
        _check(lib().msbfs_upload_queries(self._h, offsets.size - 1, _ptr(offsets), _ptr(sources) if sources.size else None))
        self._k = offsets.size - 1

    def run_resident(self) -> None:
        _check(lib().msbfs_run_resident(self._h))

    def download_sums(self, count: int | None = None) -> np.ndarray:
        k = self._k if count is None else count
        sums = np.empty(max(k, 1), dtype=np.int64)
        _check(lib().msbfs_download_sums(self._h, _ptr(sums), k))
        return sums[:k]

    def reached_entries(self, count: int | None = None) -> np.ndarray:
        k = self._k if count is None else count
        out = np.empty(max(k, 1), dtype=np.int64)
        _check(lib().msbfs_reached_entries(self._h, _ptr(out), k))
        return out[:k]

    def vertex_degrees(self, vertex_ids: np.ndarray) -> np.ndarray:
        ids = np.ascontiguousarray(vertex_ids, dtype=np.int32)
        out = np.zeros(max(ids.size, 1), dtype=np.int64)
        if ids.size:
            _check(lib().msbfs_vertex_degrees(self._h, ids.size, _ptr(ids), _ptr(out)))
        return out[: ids.size]

    def group_weights(self, offsets: np.ndarray, sources: np.ndarray) -> np.ndarray:
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        sources = np.ascontiguousarray(sources, dtype=np.int32)
        k = offsets.size - 1
        out = np.zeros(max(k, 1), dtype=np.int64)
        if k:
            _check(lib().msbfs_group_weights(self._h, k, _ptr(offsets), _ptr(sources) if sources.size else None, _ptr(out)))
        return out[:k]

    def stats(self) -> dict:
        st = Stats()
        _check(lib().msbfs_get_stats(self._h, C.byref(st)))
        return st.as_dict()

    def level_log(self, capacity: int = 1 << 16) -> list:
        buf = (LevelRecord * capacity)()
        cnt = C.c_int32()
        _check(lib().msbfs_get_level_log(self._h, buf, capacity, C.byref(cnt)))
        return [{name: getattr(buf[i], name) for name, _ in LevelRecord._fields_} for i in range(min(cnt.value, capacity))]

    def launch_log(self, capacity: int = 1 << 16) -> list:
        buf = (LaunchRecord * capacity)()
        cnt = C.c_int32()
        _check(lib().msbfs_get_launch_log(self._h, buf, capacity, C.byref(cnt)))
        return [{"batch": buf[i].batch, "level": buf[i].level, "slot": buf[i].slot, "ms": buf[i].ms}
                for i in range(min(cnt.value, capacity))]

    def batch_ms(self, capacity: int = 4096) -> list:
        buf = (C.c_double * capacity)()
        cnt = C.c_int32()
        _check(lib().msbfs_get_batch_ms(self._h, buf, capacity, C.byref(cnt)))
        return [buf[i] for i in range(min(cnt.value, capacity))]

    def kernel_profiles(self) -> list:
        out = []
        for slot in range(lib().msbfs_num_kernel_slots()):
            kp = KernelProfile()
            _check(lib().msbfs_get_kernel_profile(self._h, slot, C.byref(kp)))
            out.append(kp.as_dict())
        return out

    def solve(self, offsets: np.ndarray, sources: np.ndarray) -> Report:
        sums = self.run(offsets, sources)
        k, f = argmin(sums)
        return Report(k, f, sums)
