"""The C++ facade (include/El.hpp) on a GPU: code written against the reference's names —
El::Grid, El::DistMatrix<T,U,V,ELEMENT,Device::GPU>, El::Matrix<T,Device::GPU>, El::Gemm,
El::LocalGemm, El::Copy / Transpose / Axpy / Scale / Zero / AxpyContract / Contract, SyncInfo,
the blocksize stack — compiled by `make` (tools/el_api_check, tools/el_gemm_driver) and run here.
el_gemm_driver is the counterpart of the reference's tests/blas_like/Gemm.cpp (orientations x
SUMMA variants, associativity residual asserted)."""
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args):
    exe = os.path.join(ROOT, "tools", args[0])
    if not os.path.exists(exe):
        pytest.fail(f"{exe} is missing: run `make` (or __graft_entry__.build()) first")
    return subprocess.run([exe] + args[1:], capture_output=True, text=True, timeout=300, cwd=ROOT)


def test_el_api_check():
    r = _run(["el_api_check"])
    assert r.returncode == 0 and "EL_API_CHECK_PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("extra", [[], ["--f64"]])
def test_el_gemm_driver(extra):
    r = _run(["el_gemm_driver", "--m", "300", "--n", "200", "--k", "250", "--reps", "1"] + extra)
    assert r.returncode == 0 and "EL_GEMM_DRIVER_PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_el_comm_check_single_rank():
    """Grid(mpi::COMM_WORLD, r), grid communicators, mpi:: collectives, proxies, MultiSync, SyncInfoPool on a 1 x 1 grid."""
    r = _run(["el_comm_check"])
    assert r.returncode == 0 and "EL_COMM_CHECK_PASS 1x1" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("nproc,height", [(2, 1), (4, 2), (8, 2)])
def test_el_comm_check_multi_gpu(nproc, height):
    """The same driver with one process per GPU under tools/elrun (the stand-in for `mpiexec -np N`): the world communicator
    bootstraps over TCP, the grid's NCCL communicators from an id broadcast over it, collectives over the peer arena."""
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    exe = os.path.join(ROOT, "tools", "el_comm_check")
    env = dict(os.environ, MASTER_PORT=str(24000 + nproc * 10 + height))
    r = subprocess.run([os.path.join(ROOT, "tools", "elrun"), str(nproc), exe, "--gridHeight", str(height)], capture_output=True, text=True,
                       timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and "EL_COMM_CHECK_PASS" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
