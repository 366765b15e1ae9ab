"""The drop-in executable: usage / missing-file paths (no GPU needed) and,
on a GPU box, the full seven-line report against the reference's golden output."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN_CASES, golden_case_inputs


def _paths(pkg):
    host = os.path.join(pkg.PKG_DIR, "host")
    return os.path.join(host, "main"), os.path.join(host, "mpirun")


def test_usage_message_and_status(pkg):
    main, mpirun = _paths(pkg)
    res = subprocess.run([main, "-g", "x"], capture_output=True, text=True)
    assert res.returncode == 255 and res.stdout == ""  # return -1 (main.cu:211)
    assert res.stderr.strip() == f"Usage: mpirun -np <ranks> {main} -g <graph.bin> -q <query.bin> -gn <numGPU>"
    res = subprocess.run([mpirun, "-np", "3", main], capture_output=True, text=True)
    assert res.returncode == 255 and res.stderr.count("Usage:") == 1  # rank 0 only (main.cu:205)


def test_missing_files(pkg, tmp_path, golden_dir):
    main, mpirun = _paths(pkg)
    res = subprocess.run([main, "-g", "nope.bin", "-q", "q.bin", "-gn", "1"], capture_output=True, text=True)
    assert res.returncode == 1 and res.stderr.strip() == "Could not open graph file nope.bin" and res.stdout == ""
    g = os.path.join(golden_dir, "kat_g7.graph.bin")
    res = subprocess.run([mpirun, "-np", "2", main, "-g", g, "-q", "nope.bin", "-gn", "1"], capture_output=True, text=True, timeout=60)
    assert res.returncode == 1 and res.stderr.strip() == "Could not open query file nope.bin"


def test_reference_host_build_same_error_paths(pkg, ob):
    """The reference's own binary under the bundled launcher: identical messages / codes."""
    if not os.path.exists(ob.REF_HOST):
        pytest.skip("host build of the reference not present")
    _, mpirun = _paths(pkg)
    res = subprocess.run([ob.REF_HOST, "-g", "x"], capture_output=True, text=True)
    assert res.returncode == 255 and res.stderr.startswith("Usage: mpirun -np <ranks> ")
    res = subprocess.run([mpirun, "-np", "2", ob.REF_HOST, "-g", "nope.bin", "-q", "q", "-gn", "1"], capture_output=True, text=True, timeout=60)
    assert res.returncode == 1 and res.stderr.strip() == "Could not open graph file nope.bin"


REPORT = re.compile(
    r"^Graph: (?P<g>.*)\nQuery: (?P<q>.*)\nQuery number \(k\) with minimum F value: (?P<k>-?\d+)\n"
    r"Minimum F value: (?P<f>-?\d+)\nGPU # : (?P<gn>\d+) GPU\nPreprocessing time: \d+\.\d{9} s\n"
    r"Computation time: \d+\.\d{9} s\n$")


@pytest.mark.gpu
@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("ranks", [1, 2])
def test_report_matches_reference(pkg, ob, expected, tmp_path, name, ranks):
    """`mpirun -np R ./main -g G -q Q -gn 1` prints the reference's seven lines with the
    reference's k and F (np 2 on one GPU = BASELINE config 1's launch shape)."""
    main, mpirun = _paths(pkg)
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    gpath = str(tmp_path / "g.bin")
    qpath = os.path.join(os.path.dirname(__file__), "golden", expected[name]["query_file"])
    pkg.write_graph_file(gpath, n, uv)
    dump = str(tmp_path / "sums.txt")
    res = subprocess.run([mpirun, "-np", str(ranks), main, "-g", gpath, "-q", qpath, "-gn", "1", "--dump-sums", dump],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    m = REPORT.match(res.stdout)
    assert m, res.stdout
    assert (m["g"], m["q"], m["gn"]) == (gpath, qpath, "1")
    assert int(m["k"]) == expected[name]["k_printed"] and int(m["f"]) == expected[name]["min_f"]
    assert res.stdout.splitlines()[2:5] == expected[name]["report_lines_3_to_5"]
    got = [int(line.split()[1]) for line in open(dump)] if os.path.exists(dump) else []
    assert got == expected[name]["sums"]


@pytest.mark.gpu
def test_ranks_on_one_gpu_share_one_csr(pkg, ob, expected, tmp_path):
    """`mpirun -np 3 ... -gn 1` (ranks > GPUs, main.cu:227): rank 0 builds the CSR, ranks 1 and 2
    map it through CUDA IPC (msbfs_export_graph / msbfs_import_graph) instead of uploading
    their own copies as the reference does (main.cu:283-291); MSBFS_NO_IPC=1 keeps one copy per
    rank.  Same report either way."""
    main, mpirun = _paths(pkg)
    name = "rmat16_64x4"
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    gpath = str(tmp_path / "g.bin")
    qpath = os.path.join(os.path.dirname(__file__), "golden", expected[name]["query_file"])
    pkg.write_graph_file(gpath, n, uv)
    outs = {}
    for no_ipc in (False, True):
        env = dict(os.environ, MSBFS_STATS="1")
        if no_ipc:
            env["MSBFS_NO_IPC"] = "1"
        res = subprocess.run([mpirun, "-np", "3", main, "-g", gpath, "-q", qpath, "-gn", "1"], capture_output=True, text=True,
                             timeout=300, env=env)
        assert res.returncode == 0, res.stderr
        m = REPORT.match(res.stdout)
        assert m, res.stdout
        assert int(m["k"]) == expected[name]["k_printed"] and int(m["f"]) == expected[name]["min_f"]
        outs[no_ipc] = res.stderr
    shared = [ln for ln in outs[False].splitlines() if "graph=" in ln]
    assert len(shared) == 3 and sum("graph=shared" in ln for ln in shared) == 2, outs[False]
    assert all("graph=own" in ln for ln in outs[True].splitlines() if "graph=" in ln)


@pytest.mark.gpu
def test_nccl_reduce_in_the_drop_in_two_gpus(pkg, ob, expected, tmp_path):
    """`mpirun -np 2 ... -gn 2`: the packed-min reduction runs over NCCL (communicator made from
    an ncclUniqueId broadcast over MPI) and the graph file is read once per box (sharded read +
    ncclAllGather of the edge records); MSBFS_NO_SHARDED_LOAD=1 / MSBFS_NO_NCCL=1 fall back to a
    whole-file read per rank / the MPI reduction.  Same seven report lines every way.  Needs two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    main, mpirun = _paths(pkg)
    name = "kron12_200x3"
    n, uv, offs, src = golden_case_inputs(pkg, ob, expected, name)
    gpath = str(tmp_path / "g.bin")
    qpath = os.path.join(os.path.dirname(__file__), "golden", expected[name]["query_file"])
    pkg.write_graph_file(gpath, n, uv)
    for mode in ("nccl", "nccl-whole-file", "mpi"):
        env = dict(os.environ, MSBFS_STATS="1", NCCL_DEBUG="VERSION")  # NCCL's banner must not reach the report
        if mode == "mpi":
            env["MSBFS_NO_NCCL"] = "1"
        if mode == "nccl-whole-file":
            env["MSBFS_NO_SHARDED_LOAD"] = "1"
        res = subprocess.run([mpirun, "-np", "2", main, "-g", gpath, "-q", qpath, "-gn", "2"], capture_output=True, text=True,
                             timeout=300, env=env)
        assert res.returncode == 0, res.stderr
        m = REPORT.match(res.stdout)
        assert m and int(m["k"]) == expected[name]["k_printed"] and int(m["f"]) == expected[name]["min_f"], res.stdout
        want = "reduce=mpi" if mode == "mpi" else "reduce=nccl"
        assert sum(want in ln for ln in res.stderr.splitlines()) == 2, res.stderr
        # with a communicator each rank reads half of the file and an all-gather completes the edge list
        load = "load=sharded" if mode == "nccl" else "load=whole"
        assert sum(load in ln for ln in res.stderr.splitlines()) == 2, res.stderr
