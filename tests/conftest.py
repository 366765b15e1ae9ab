import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as graft

    p = graft.load_package()
    if not os.path.exists(p.LIB_PATH):
        p.build()
    return p


@pytest.fixture(scope="session")
def ob():
    import oracle_binding

    oracle_binding.lib()
    return oracle_binding


@pytest.fixture(scope="session")
def expected():
    with open(os.path.join(GOLDEN, "expected.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def golden_case_inputs(pkg, ob, expected, name):
    """(n, uv, offsets, sources) of a golden case: explicit files for the KATs,
    regenerated graph + committed query file for the generated ones."""
    case = expected[name]
    offs, src = ob.load_query_file(os.path.join(GOLDEN, case["query_file"]))
    if "graph_file" in case:
        n, uv = ob.load_graph_file(os.path.join(GOLDEN, case["graph_file"]))
    else:
        g = case["gen"]
        n = pkg.gen_num_vertices(g["kind"], g["p0"], g["p1"])
        uv = pkg.gen_edges_host(g["kind"], g["p0"], g["p1"], g["m"], g["seed"])
    return n, uv, offs, src


GOLDEN_CASES = ["kat1", "kat2", "kat3", "kat4", "rmat16_64x4", "rmat12_ragged", "kron12_200x3", "lattice_48x40", "er_4096"]
