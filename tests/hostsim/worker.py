"""Runs tests/dist_cases.py on an r x c grid of THREAD ranks against the hostsim build of the
engine (no GPU):  python tests/hostsim/worker.py RxC [--heavy] [--only name,...]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("grid")
    ap.add_argument("--heavy", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--level3", action="store_true", help="run dist_cases.run_level3 instead of run_all")
    args = ap.parse_args()
    grids = [tuple(int(x) for x in gs.lower().split("x")) for gs in args.grid.split(",")]   # "2x2" or "1x2,2x1"
    from tests.hostsim import sim_el
    from tests import dist_cases as dc
    dc.use_backend(sim_el)
    only = [x for x in args.only.split(",") if x]

    def body(grid):
        if args.level3:
            return dc.run_level3(grid)
        if not only:
            return dc.run_all(grid, heavy=args.heavy)
        out = {}
        import torch
        for name in only:
            # check_xxx_f32 / check_xxx_f64 -> dc.check_xxx(grid, dtype); bare names take the case's defaults
            if name.endswith("_f32") or name.endswith("_f64"):
                out[name] = getattr(dc, name[:-4])(grid, torch.float32 if name.endswith("_f32") else torch.float64)
            else:
                out[name] = getattr(dc, name)(grid)
        return out

    for (r, c) in grids:
        t0 = time.time()
        res = sim_el.run_grid(r, c, body)
        print(f"hostsim grid {r}x{c}: {res[0]} in {time.time() - t0:.1f} s", flush=True)
    print("HOSTSIM_PASS", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
