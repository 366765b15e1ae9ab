"""The bundled single-node MPI subset (host/minimpi) on its own: every collective the host
program and the unmodified reference call, payloads larger than the shared window, ranks
that contribute nothing, and the launcher's exit-status rules.  CPU only."""
import os
import subprocess

import pytest


@pytest.fixture(scope="module")
def check_exe(pkg, tmp_path_factory):
    mpidir = os.path.join(pkg.PKG_DIR, "host", "minimpi")
    exe = str(tmp_path_factory.mktemp("minimpi") / "minimpi_check")
    src = os.path.join(os.path.dirname(__file__), "native", "minimpi_check.c")
    subprocess.run(["gcc", "-O1", "-std=c11", "-I", mpidir, src, os.path.join(mpidir, "minimpi.c"), "-o", exe, "-lrt"], check=True)
    return exe


@pytest.mark.parametrize("ranks", [1, 2, 3, 7])
def test_collectives(pkg, check_exe, ranks):
    mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
    env = dict(os.environ, MINIMPI_WINDOW_MB="1")  # force multi-chunk broadcasts
    res = subprocess.run([mpirun, "-np", str(ranks), check_exe], capture_output=True, text=True, timeout=120, env=env)
    assert res.returncode == 0, res.stderr
    assert res.stdout.strip() == f"minimpi OK with {ranks} ranks"


def test_singleton_without_launcher(check_exe):
    res = subprocess.run([check_exe], capture_output=True, text=True, timeout=60)
    assert res.returncode == 0 and "minimpi OK with 1 ranks" in res.stdout


def test_launcher_reports_failing_rank(pkg, tmp_path):
    """A rank that dies while its peers wait in a collective ends the job with its status."""
    mpirun = os.path.join(pkg.PKG_DIR, "host", "mpirun")
    mpidir = os.path.join(pkg.PKG_DIR, "host", "minimpi")
    src = tmp_path / "die.c"
    src.write_text('#include <mpi.h>\n#include <stdlib.h>\nint main(int c,char**v){int r;MPI_Init(&c,&v);MPI_Comm_rank(MPI_COMM_WORLD,&r);'
                   'if(r==1)exit(9);MPI_Barrier(MPI_COMM_WORLD);MPI_Finalize();return 0;}\n')
    exe = str(tmp_path / "die")
    subprocess.run(["gcc", "-O1", "-I", mpidir, str(src), os.path.join(mpidir, "minimpi.c"), "-o", exe, "-lrt"], check=True)
    res = subprocess.run([mpirun, "-np", "3", exe], capture_output=True, text=True, timeout=60)
    assert res.returncode == 9
