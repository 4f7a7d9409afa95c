#!/usr/bin/env bash
# One GPU session that validates and measures everything written without a GPU.
#   1 GPU : /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_session.sh 1'
#   N GPUs: /usr/local/graft/bin/gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_session.sh N'
# Everything lands in gpurun_out/ (copy what should be judged into profiles/). Nothing here runs under a
# profiler except the explicit ncu lines at the end, which come after the plain runs have exited 0.
set -u
N=${1:-1}
OUT=gpurun_out
mkdir -p $OUT
run() { echo "=== $*" | tee -a $OUT/session.log; "$@" >> $OUT/session.log 2>&1; echo "exit $?" | tee -a $OUT/session.log; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"

if [ "$N" = "1" ]; then
  # 1. parity first: the Gemm-path files, then the level-3 family (new kernels: trapezoid, trsm)
  run timeout 1200 python -m pytest tests -m gpu -x -q
  # 2. the bench lines (fp32 = BASELINE C2; fp64); e2e now has four block slots + streamed A
  python bench.py --steps 5 --warmup 3 > $OUT/bench_f32_1gpu.json 2>> $OUT/session.log
  python bench.py --steps 5 --warmup 3 --dtype f64 > $OUT/bench_f64_1gpu.json 2>> $OUT/session.log
  # 2b. host-operand Gemm A/B: uniform blocks (the measured form), tapered tail (default), taper + B in 4 k ranges
  for v in "ELB200_HOST_TAPER=0" "ELB200_HOST_TAPER=1" "ELB200_HOST_BRANGES=4" "ELB200_HOST_BRANGES=8"; do
    echo "=== e2e $v" >> $OUT/e2e_variants.log; env $v timeout 120 tools/native_bench e2e >> $OUT/e2e_variants.log 2>&1
  done
  ELB200_HOST_BRANGES=4 timeout 600 python -m pytest tests/test_gpu_dist.py tests/test_gpu_level3.py -m gpu -x -q -k "gemm_host" > $OUT/e2e_branges_pytest.log 2>&1; echo "BRANGES pytest exit $?" >> $OUT/session.log
  # 3. fp64 kernel A/B: current vs warp-specialised (ELB200_F64_WS=1), same binary; the kernel tests with the variant on
  ELB200_F64_WS=1 timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k gemm > $OUT/f64_ws_pytest.log 2>&1; echo "f64 WS pytest exit $?" >> $OUT/session.log
  run tools/gemm_check perf f64 8192 16384
  ELB200_F64_WS=1 tools/gemm_check correctness f64 > $OUT/f64_ws_correctness.log 2>&1; echo "f64 WS correctness exit $?" >> $OUT/session.log
  ELB200_F64_WS=1 tools/gemm_check perf f64 8192 16384 > $OUT/f64_ws_perf.log 2>&1
  # 3b. TMA-fed fp64 kernel (experimental, never run before): correctness first, then time
  ELB200_F64_TMA=1 timeout 60 tools/native_bench f64 8192 > $OUT/f64_tma_native.log 2>&1; echo "f64 TMA native exit $?" >> $OUT/session.log
  ELB200_F64_TMA=1 timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k gemm > $OUT/f64_tma_pytest.log 2>&1; echo "f64 TMA pytest exit $?" >> $OUT/session.log
  # 4. new HBM-bound kernels and the level-3 family
  run python tools/blas1_bw.py --out $OUT/blas1_bandwidth.json
  run python tools/level3_bench.py --n 16384 --out $OUT/level3_1gpu.json
  # 5. launch list + one full capture of the trapezoid kernel (after the plain runs above)
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_level3.csv \
      python tools/level3_bench.py --n 8192 --dtype f32 > $OUT/ncu_level3.log 2>&1
else
  run timeout 900 $TR tests/dist_worker.py
  run timeout 900 $TR tests/dist_worker.py --level3
  $TR bench.py --gpus $N --steps 5 --warmup 3 > $OUT/bench_f32_${N}gpu.json 2>> $OUT/session.log
  $TR bench.py --gpus $N --steps 5 --warmup 3 --dtype f64 > $OUT/bench_f64_${N}gpu.json 2>> $OUT/session.log
  # C4 with the split AllGather, and with it off; C5 with the SUMMA-Dot pipeline, and with it off
  # the split gather is opt-in until it has passed here: parity with it on, then the A/B
  ELB200_AG_SPLIT_MB=4 timeout 900 $TR tests/dist_worker.py > $OUT/dist_worker_agsplit.log 2>&1; echo "AG split dist_worker exit $?" >> $OUT/session.log
  ELB200_DOT_PIPELINE=0 $TR tools/dist_sweep.py --c4 --c5 --out $OUT/dist_sweep_${N}gpu_before.json >> $OUT/session.log 2>&1
  ELB200_AG_SPLIT_MB=4 timeout 900 $TR tools/dist_sweep.py --c4 --c5 --out $OUT/dist_sweep_${N}gpu.json >> $OUT/session.log 2>&1
  ELB200_AG_SPLIT_MB=4 ELB200_AG_ROTATE=0 timeout 900 $TR tools/dist_sweep.py --c4 --out $OUT/dist_sweep_${N}gpu_norotate.json >> $OUT/session.log 2>&1
  run $TR tools/level3_bench.py --n 32768 --dtype f32 --out $OUT/level3_${N}gpu.json
fi
tail -5 $OUT/session.log
