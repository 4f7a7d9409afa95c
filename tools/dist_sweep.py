"""SURVEY.md 8(d) configs C4 (redistribution sweep) and C5 (SUMMA_Dot, tall-k) on libelb200.

Launch exactly like bench.py: `python tools/dist_sweep.py --c4 --c5` for one GPU, or
`python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
--master-port P tools/dist_sweep.py --c4 --c5 --out profiles/r01_dist_sweep_Ngpu.json`.

C4: for global sizes S, square fp32/fp64 matrices, time on the grid's stream
      [MC,MR] -> [MC,*]   (copy::RowAllGather,       Copy/RowAllGather.hpp:17-175)
      [MC,MR] -> [*,VR]   (copy::RowAllToAllDemote,  Copy/RowAllToAllDemote.hpp)
      AxpyContract([MC,*] -> [MC,MR])  (axpy_contract::RowScatter, AxpyContract.hpp:146-260)
    and report S / t (the reference's own "redistribution GB/s") plus the per-GPU NVLink bytes
    (g-1)/g * local output (allgather / all-to-all) or (g-1)/g * local input (reduce-scatter) / t.
C5: fp32 NN m = n = 8192, k = 262144, GEMM_DEFAULT (must pick SUMMA_Dot, Gemm/NN.hpp:592-593).
Times are CUDA events on the grid stream, max over ranks; never run under a profiler.
"""
import argparse
import gc
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(grid, fn, world, local, warm=2, reps=5):
    for _ in range(warm):
        fn()
    grid.sync()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(grid.stream)
    for _ in range(reps):
        fn()
    e1.record(grid.stream)
    grid.sync()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=f"cuda:{local}", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def fill(grid, M, seed):
    g = torch.Generator(device=M.Matrix().device).manual_seed(seed)
    with torch.cuda.stream(grid.stream):
        v = M.Matrix()
        v.copy_((torch.rand(v.shape, device=v.device, dtype=v.dtype, generator=g) - 0.5) * 0.2)


def verify_gather(grid, A, B1, world, local):
    """[MC,MR] -> [MC,*] moved every entry exactly once: the sum of my gathered local matrix equals the sum of the local
    sums of the ranks of my grid row (all entries are exactly representable partial sums in fp64 up to rounding)."""
    dev = f"cuda:{local}"
    with torch.cuda.stream(grid.stream):
        mine = A.Matrix().double().sum().reshape(1)
        got = B1.Matrix().double().sum()
    grid.sync()
    sums = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
    if world > 1:
        dist.all_gather(sums, mine)
    else:
        sums = [mine]
    r = grid.Height()
    want = sum(float(sums[q].item()) for q in range(world) if q % r == grid.MCRank())
    scale = max(abs(want), 1.0)
    return abs(float(got.item()) - want) / scale


def run_c4(el, grid, world, rank, local, max_bytes, rows, min_bytes=0):
    r, c = grid.Height(), grid.Width()
    for tdt, name, esz in ((torch.float32, "f32", 4), (torch.float64, "f64", 8)):
        for S in (1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30, 4 << 30, 8 << 30):
            if S > max_bytes or S < min_bytes:
                continue
            n = int((S / esz) ** 0.5) // 64 * 64
            A = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(n, n)
            fill(grid, A, 5 + rank)
            B1 = el.DistMatrix(grid, el.MC, el.STAR, tdt).Resize(n, n)
            t1 = timed(grid, lambda: el.Copy(A, B1), world, local)
            out1 = B1.LocalHeight() * B1.LocalWidth() * esz
            if S >= (1 << 30):
                err = verify_gather(grid, A, B1, world, local)
                assert err < 1e-9, f"gather checksum off by {err:.3e} at S={S}"
            B2 = el.DistMatrix(grid, el.STAR, el.VR, tdt).Resize(n, n)
            t2 = timed(grid, lambda: el.Copy(A, B2), world, local)
            out2 = B2.LocalHeight() * B2.LocalWidth() * esz
            t3 = timed(grid, lambda: el.AxpyContract(1e-3, B1, A), world, local)
            glob = n * n * esz
            p = r * c
            for op, t, g, loc in (("[MC,MR]->[MC,*] RowAllGather", t1, c, out1),
                                  ("[MC,MR]->[*,VR] RowAllToAllDemote", t2, p, out2),
                                  ("AxpyContract [MC,*]->[MC,MR] RowScatter", t3, c, out1)):
                link = (g - 1) / g * loc
                rows.append({"config": "C4", "op": op, "dtype": name, "grid": f"{r}x{c}", "n": n, "global_bytes": glob,
                             "ms": t, "global_GBps": glob / t / 1e6, "nvlink_bytes_per_gpu": link,
                             "nvlink_GBps_per_gpu": link / t / 1e6})
                if rank == 0:
                    print(f"C4 {name} {r}x{c} n={n:6d} S={glob / 2**20:8.1f} MiB  {op:42s} {t:9.3f} ms  "
                          f"{glob / t / 1e6:8.1f} GB/s global  {link / t / 1e6:7.1f} GB/s/GPU link", flush=True)
            del A, B1, B2
            gc.collect()
            grid.sync()


def run_c5(el, grid, world, rank, local, k, rows):
    m = n = 8192
    r, c = grid.Height(), grid.Width()
    tdt = torch.float32
    A = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(m, k)
    B = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(k, n)
    Cm = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(m, n)
    for i, M in enumerate((A, B, Cm)):
        fill(grid, M, 11 * (i + 1) + rank)
    for alg, label in ((el.GEMM_DEFAULT, "GEMM_DEFAULT (-> SUMMA_Dot)"), (el.GEMM_SUMMA_C, "GEMM_SUMMA_C")):
        t = timed(grid, lambda: el.Gemm(el.NORMAL, el.NORMAL, 0.5, A, B, -0.5, Cm, alg), world, local, warm=1, reps=3)
        tf = 2.0 * m * n * k / t / 1e9
        rows.append({"config": "C5", "alg": label, "dtype": "f32", "grid": f"{r}x{c}", "m": m, "n": n, "k": k, "ms": t,
                     "TFLOPs": tf, "TFLOPs_per_gpu": tf / (r * c)})
        if rank == 0:
            print(f"C5 f32 {r}x{c} m=n={m} k={k} {label:30s} {t:9.2f} ms  {tf:8.1f} TFLOP/s  ({tf / (r * c):.1f} per GPU)", flush=True)
    del A, B, Cm
    gc.collect()
    grid.sync()


def run_c6(el, grid, world, rank, local, rows):
    """Stationary-A / stationary-B SUMMA on shapes with several panels (not a BASELINE config; the A/B shape for the
    SUMMA-A / B panel pipeline, ELB200_AB_PIPELINE): fp32, m = k = 32768, n = 16384 for A; m = 16384, n = k = 32768 for B."""
    r, c = grid.Height(), grid.Width()
    tdt = torch.float32
    for alg, label, (m, n, k) in ((el.GEMM_SUMMA_A, "GEMM_SUMMA_A", (32768, 16384, 32768)), (el.GEMM_SUMMA_B, "GEMM_SUMMA_B", (16384, 32768, 32768))):
        A = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(m, k)
        B = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(k, n)
        Cm = el.DistMatrix(grid, el.MC, el.MR, tdt).Resize(m, n)
        for i, M in enumerate((A, B, Cm)):
            fill(grid, M, 17 * (i + 1) + rank)
        t = timed(grid, lambda: el.Gemm(el.NORMAL, el.NORMAL, 0.5, A, B, -0.5, Cm, alg), world, local, warm=1, reps=3)
        tf = 2.0 * m * n * k / t / 1e9
        rows.append({"config": "C6", "alg": label, "dtype": "f32", "grid": f"{r}x{c}", "m": m, "n": n, "k": k, "ms": t, "TFLOPs": tf,
                     "TFLOPs_per_gpu": tf / (r * c), "ab_pipeline": os.environ.get("ELB200_AB_PIPELINE", "default")})
        if rank == 0:
            print(f"C6 f32 {r}x{c} m={m} n={n} k={k} {label:14s} {t:9.2f} ms  {tf:8.1f} TFLOP/s  ({tf / (r * c):.1f} per GPU)", flush=True)
        del A, B, Cm
        gc.collect()
        grid.sync()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--c4", action="store_true")
    ap.add_argument("--c5", action="store_true")
    ap.add_argument("--c6", action="store_true", help="SUMMA-A / SUMMA-B on multi-panel shapes")
    ap.add_argument("--max-bytes", type=int, default=1 << 30, help="largest global matrix of the C4 sweep")
    ap.add_argument("--min-bytes", type=int, default=0, help="smallest global matrix of the C4 sweep")
    ap.add_argument("--k", type=int, default=262144)
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
    rows = []
    try:
        if args.c4:
            run_c4(el, grid, world, rank, local, args.max_bytes, rows, args.min_bytes)
        if args.c5:
            run_c5(el, grid, world, rank, local, args.k, rows)
        if args.c6:
            run_c6(el, grid, world, rank, local, rows)
    finally:
        gc.collect()
        torch.cuda.synchronize()
        grid.close()
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump({"n_gpus": world, "rows": rows}, f, indent=1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
