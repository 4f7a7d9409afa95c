"""torchrun worker: the distributed parity cases on an r x c grid, one process per GPU (NCCL).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P tests/dist_worker.py [--height r] [--heavy] [--level3] [--only check_complex,...]
"""
import argparse
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--heavy", action="store_true")
    ap.add_argument("--level3", action="store_true", help="run dist_cases.run_level3 instead of run_all")
    ap.add_argument("--only", default="", help="comma-separated dist_cases functions taking (grid) to run instead of run_all")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from elemental_b200 import el
    from tests import dist_cases as dc
    ok = 1
    try:
        grid = el.Grid(height=args.height or None, device=local)
        if args.only:
            out = {name: getattr(dc, name)(grid) for name in args.only.split(",")}
        else:
            out = dc.run_level3(grid) if args.level3 else dc.run_all(grid, heavy=args.heavy)
        grid.sync()
        if grid.Rank() == 0:
            print(f"grid {grid.Height()}x{grid.Width()}: {out}", flush=True)
    except Exception:
        traceback.print_exc()
        ok = 0
    flag = torch.tensor([ok], device=f"cuda:{local}")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if dist.get_rank() == 0:
        print("DIST_WORKER_PASS" if int(flag.item()) else "DIST_WORKER_FAIL", flush=True)
    dist.destroy_process_group()
    return 0 if int(flag.item()) else 1


if __name__ == "__main__":
    sys.exit(main())
