"""TEST INFRASTRUCTURE: ctypes binding of oracle/liboracle.so (the CPU restatement of
/root/reference/main.cu).  Imported only by tests, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs — never by the product path."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "liboracle.so")
REF_HOST = os.path.join(ORACLE_DIR, "_ref", "main_host")
REF_GPU = os.path.join(ORACLE_DIR, "_ref", "main_gpu")

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, capture_output=True)
        h = C.CDLL(LIB)
        h.oracle_build_csr.restype = C.c_int
        h.oracle_build_csr.argtypes = [C.c_int32, C.c_int64, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        h.oracle_build_csr_parallel.restype = C.c_int
        h.oracle_build_csr_parallel.argtypes = h.oracle_build_csr.argtypes
        h.oracle_load_graph_bin.restype = C.c_int
        h.oracle_load_graph_bin.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_void_p)]
        h.oracle_load_query_bin.restype = C.c_int
        h.oracle_load_query_bin.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        h.oracle_group_distances.restype = C.c_int
        h.oracle_group_distances.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        h.oracle_run.restype = C.c_int
        h.oracle_run.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
        h.oracle_run_wide.restype = C.c_int
        h.oracle_run_wide.argtypes = h.oracle_run.argtypes
        h.oracle_set_threads.restype = None
        h.oracle_set_threads.argtypes = [C.c_int]
        h.oracle_argmin.restype = None
        h.oracle_argmin.argtypes = [C.c_int32, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int64)]
        h.oracle_num_threads.restype = C.c_int
        h.oracle_free.restype = None
        h.oracle_free.argtypes = [C.c_void_p]
        _lib = h
    return _lib


def _p(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


class Csr:
    """CSR built by the oracle (record order per row, both directions; main.cu:106-129)."""

    def __init__(self, n: int, uv: np.ndarray, parallel: bool = False):
        uv = np.ascontiguousarray(uv, dtype=np.int32).reshape(-1, 2)
        self.n, self.m = int(n), int(uv.shape[0])
        row, col = C.c_void_p(), C.c_void_p()
        build = lib().oracle_build_csr_parallel if parallel else lib().oracle_build_csr
        rc = build(self.n, self.m, _p(uv) if uv.size else None, C.byref(row), C.byref(col))
        if rc != 0:
            raise ValueError(f"oracle_build_csr failed with {rc}")
        self._row, self._col = row, col
        self.row = np.ctypeslib.as_array(C.cast(row, C.POINTER(C.c_int64)), shape=(self.n + 1,))
        self.col = np.ctypeslib.as_array(C.cast(col, C.POINTER(C.c_int32)), shape=(max(2 * self.m, 1),))[: 2 * self.m]

    def __del__(self):
        try:
            lib().oracle_free(self._row)
            lib().oracle_free(self._col)
        except Exception:
            pass


@dataclass
class Result:
    sums: np.ndarray
    reached_entries: np.ndarray
    min_k: int
    min_f: int


def argmin(sums: np.ndarray):
    sums = np.ascontiguousarray(sums, dtype=np.int64)
    k, f = C.c_int32(), C.c_int64()
    lib().oracle_argmin(sums.size, _p(sums) if sums.size else None, C.byref(k), C.byref(f))
    return k.value, f.value


def host_threads() -> int:
    """Cores this process may run on (not OMP_NUM_THREADS, which launchers set to 1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return os.cpu_count() or 1


def set_threads(n: int) -> None:
    lib().oracle_set_threads(int(n))


def run_csr(csr: Csr, offsets: np.ndarray, sources: np.ndarray, threads: int = 0, wide: bool = False) -> Result:
    """F of every group.  wide=False: one BFS per thread; wide=True: groups in turn, all
    threads inside each BFS (same results)."""
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    sources = np.ascontiguousarray(sources, dtype=np.int32)
    if sources.size == 0:
        sources = np.zeros(1, dtype=np.int32)
    k = offsets.size - 1
    sums = np.zeros(max(k, 1), dtype=np.int64)
    reached = np.zeros(max(k, 1), dtype=np.int64)
    fn = lib().oracle_run_wide if wide else lib().oracle_run
    rc = fn(csr.n, C.c_void_p(csr._row.value), C.c_void_p(csr._col.value), k, _p(offsets), _p(sources),
                          _p(sums), _p(reached), threads)
    if rc != 0:
        raise RuntimeError(f"oracle_run failed with {rc}")
    mk, mf = argmin(sums[:k])
    return Result(sums[:k], reached[:k], mk, mf)


def run(n: int, uv: np.ndarray, offsets: np.ndarray, sources: np.ndarray, threads: int = 0) -> Result:
    return run_csr(Csr(n, uv), offsets, sources, threads)


def distances(csr: Csr, sources: np.ndarray) -> np.ndarray:
    sources = np.ascontiguousarray(sources, dtype=np.int32)
    dist = np.empty(max(csr.n, 1), dtype=np.int32)
    lib().oracle_group_distances(csr.n, C.c_void_p(csr._row.value), C.c_void_p(csr._col.value),
                                 _p(sources) if sources.size else None, sources.size, _p(dist))
    return dist[: csr.n]


def load_graph_file(path: str):
    n, m, uv = C.c_int32(), C.c_int64(), C.c_void_p()
    rc = lib().oracle_load_graph_bin(path.encode(), C.byref(n), C.byref(m), C.byref(uv))
    if rc != 0:
        raise OSError(f"oracle_load_graph_bin({path}) failed with {rc}")
    arr = np.ctypeslib.as_array(C.cast(uv, C.POINTER(C.c_int32)), shape=(max(2 * m.value, 1),))[: 2 * m.value].copy()
    lib().oracle_free(uv)
    return n.value, arr.reshape(-1, 2)


def load_query_file(path: str):
    k, po, ps = C.c_int32(), C.c_void_p(), C.c_void_p()
    rc = lib().oracle_load_query_bin(path.encode(), C.byref(k), C.byref(po), C.byref(ps))
    if rc != 0:
        raise OSError(f"oracle_load_query_bin({path}) failed with {rc}")
    offs = np.ctypeslib.as_array(C.cast(po, C.POINTER(C.c_int64)), shape=(k.value + 1,)).copy()
    total = int(offs[-1])
    src = np.ctypeslib.as_array(C.cast(ps, C.POINTER(C.c_int32)), shape=(max(total, 1),))[:total].copy()
    lib().oracle_free(po)
    lib().oracle_free(ps)
    return offs, src


def parse_report(stdout: str):
    """(k_printed, F) from the reference's seven report lines (main.cu:407-413)."""
    k = f = None
    for line in stdout.splitlines():
        if line.startswith("Query number (k) with minimum F value:"):
            k = int(line.split(":")[1])
        elif line.startswith("Minimum F value:"):
            f = int(line.split(":")[1])
    if k is None or f is None:
        raise ValueError("not a report:\n" + stdout)
    return k, f


def run_reference_host(graph_path: str, query_path: str, ranks: int = 1, timeout: float = 600.0):
    """Run the reference's own source (host build) on files; returns (k_printed, F, stdout)."""
    if not os.path.exists(REF_HOST):
        raise FileNotFoundError(REF_HOST)
    cmd = [REF_HOST, "-g", graph_path, "-q", query_path, "-gn", "1"]
    if ranks > 1:
        pkg = os.path.join(ROOT, "parallel-multi-source-bfs-implementation-using-mpi-and-cuda_b200")
        cmd = [os.path.join(pkg, "host", "mpirun"), "-np", str(ranks)] + cmd
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if res.returncode != 0:
        raise RuntimeError(f"reference exited with {res.returncode}: {res.stderr}")
    k, f = parse_report(res.stdout)
    return k, f, res.stdout
