"""N>1 on CPU: real processes, torch.distributed over gloo, the planner of libelb200 called per
rank (tests/gloo_worker.py).  Covers every redistribution pair plus SUMMA-C / SUMMA-A / SUMMA-Dot
panel loops on 1x2, 2x1 (world 2) and 2x2 (world 4) grids."""
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(world: int, grid_height: int, timeout=300):
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), ELB_GRID_HEIGHT=str(grid_height),
                   OMP_NUM_THREADS="1", GLOO_SOCKET_IFNAME="lo")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "gloo_worker.py")], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT))
    outs = []
    try:
        for p in procs:
            out, _ = p.communicate(timeout=timeout)
            outs.append(out)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()  # exactly the processes started above
    for rank, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {rank} failed:\n{out[-3000:]}"
        assert "GLOO_WORKER_PASS" in out, out[-3000:]


@pytest.mark.parametrize("world,height", [(2, 1), (2, 2), (4, 2)])
def test_planner_drives_real_collectives(world, height):
    _launch(world, height)
