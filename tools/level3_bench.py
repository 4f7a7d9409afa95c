"""Throughput of the triangular rank-k family next to El::Gemm on the same operands.

Launch like bench.py: `python tools/level3_bench.py [--n 16384] [--out profiles/rNN_level3_1gpu.json]` on one GPU, or
`python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P
tools/level3_bench.py ...` on an r x c grid.  CUDA-event times on the grid stream, max over ranks, after
warm-up; useful flops: Gemm 2 n^2 k, Syrk / Trrk n^2 k (one triangle), Symm 2 n^2 m, Trsm n^2 m.
Diagonally dominant triangles for Trsm (as in the parity tests).  Never run under a profiler."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools.dist_sweep import fill, timed  # noqa: E402


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=16384)
    ap.add_argument("--k", type=int, default=0, help="inner dimension of Syrk / Trrk / Gemm (default n)")
    ap.add_argument("--m", type=int, default=0, help="right-hand sides of Symm / Trsm (default n)")
    ap.add_argument("--dtype", choices=["f32", "f64", "both"], default="both")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    from elemental_b200 import el
    grid = el.Grid(device=local)
    n, k, m = args.n, args.k or args.n, args.m or args.n
    rows = []
    for tdt, name in ((torch.float32, "f32"), (torch.float64, "f64")):
        if args.dtype not in (name, "both"):
            continue
        A = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, k)
        C = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, n)
        S = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, n)
        B = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, m)
        X = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, m)
        for i, M in enumerate((A, C, S, B, X)):
            fill(grid, M, 31 + i + 7 * rank)
        # a well conditioned triangle: |entries| <= 0.1, so n on the diagonal dominates every row
        D = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, n)
        el.Zero(D)
        with torch.cuda.stream(grid.stream):
            d = D.Matrix()
            gi = D.ColShift() + D.ColStride() * torch.arange(d.shape[0], device=d.device)
            gj = D.RowShift() + D.RowStride() * torch.arange(d.shape[1], device=d.device)
            d[gi[:, None] == gj[None, :]] = float(n)
        el.Axpy(1.0, D, S)
        grid.sync()
        cases = [
            ("Gemm NT (A A^T), SUMMA_C", 2.0 * n * n * k, lambda: el.Gemm(el.NORMAL, el.TRANSPOSE, 0.5, A, A, -0.5, C, el.GEMM_SUMMA_C)),
            ("Syrk LOWER NORMAL", 1.0 * n * n * k, lambda: el.Syrk(el.LOWER, el.NORMAL, 0.5, A, -0.5, C)),
            ("Trrk UPPER (N,T)", 1.0 * n * n * k, lambda: el.Trrk(el.UPPER, el.NORMAL, el.TRANSPOSE, 0.5, A, A, -0.5, C)),
            ("Symm LEFT LOWER", 2.0 * n * n * m, lambda: el.Symm(el.LEFT, el.LOWER, 0.5, S, B, -0.5, X)),
            ("Trsm LEFT LOWER NORMAL", 1.0 * n * n * m, lambda: (el.Copy(B, X), el.Trsm(el.LEFT, el.LOWER, el.NORMAL, el.NON_UNIT, 1.0, S, X))),
            ("Trsm RIGHT UPPER TRANSPOSE", 1.0 * n * n * m, lambda: (el.Copy(B, X), el.Trsm(el.RIGHT, el.UPPER, el.TRANSPOSE, el.NON_UNIT, 1.0, S, X)) if m == n else None),
            ("Copy (subtracted from the Trsm lines)", 0.0, lambda: el.Copy(B, X)),
            ("MakeSymmetric LOWER", 0.0, lambda: el.MakeSymmetric(el.LOWER, C)),
        ]
        copy_ms = None
        for label, flop, fn in cases:
            if "RIGHT" in label and m != n:
                continue
            ms = timed(grid, fn, world, local, warm=2, reps=3)
            if label.startswith("Copy"):
                copy_ms = ms
            rows.append({"op": label, "dtype": name, "n": n, "k": k, "m": m, "gpus": world, "grid": f"{grid.Height()}x{grid.Width()}",
                         "ms": ms, "useful_TFLOPs": (flop / (ms * 1e-3) / 1e12) if flop else None})
            if rank == 0:
                print(f"{name} {label:40s} {ms:9.3f} ms" + (f"  {flop / (ms * 1e-3) / 1e12:8.1f} TFLOP/s (useful)" if flop else ""), flush=True)
        if copy_ms is not None:
            for r in rows:
                if r["dtype"] == name and r["op"].startswith("Trsm"):
                    r["ms_minus_copy"] = r["ms"] - copy_ms
        for M in (A, C, S, B, X, D):
            M.destroy()
    grid.sync()
    if rank == 0 and args.out:
        json.dump({"note": "CUDA-event times on the grid stream, max over ranks, 3 repetitions after 2 warm-ups", "rows": rows}, open(args.out, "w"), indent=1)
    import gc
    gc.collect()
    torch.cuda.synchronize()
    grid.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
