#!/usr/bin/env bash
# Round-2 GPU sessions. Everything lands in gpurun_out/<tag>/ (copy what should be judged into profiles/).
# Plain runs first; anything under ncu comes after the same command has exited 0 without it.
#   1 GPU : gpurun --timeout 1500 -- 'bash tools/gpu_session_r2.sh 1a'
#   N GPUs: gpurun --gpus N --timeout 900 -- 'bash tools/gpu_session_r2.sh multi N'
set -u
STAGE=${1:-1a}
N=${2:-1}
OUT=gpurun_out/$STAGE${2:+_$2}
mkdir -p $OUT
log() { echo "$*" | tee -a $OUT/session.log; }
run() { log "=== $*"; "$@" >> $OUT/session.log 2>&1; log "exit $?"; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"

case $STAGE in
1a)
  # parity suite (incl. tests/test_zz_no_alloc.py, collected last), the new bench line, fp64 kernel variants
  log "=== pytest -m gpu"; timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; log "exit $?"; tail -3 $OUT/pytest_gpu.log | tee -a $OUT/session.log
  log "=== pytest no_alloc alone (fresh process)"; timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/pytest_no_alloc_alone.log 2>&1; log "exit $?"; tail -2 $OUT/pytest_no_alloc_alone.log | tee -a $OUT/session.log
  log "=== bench.py"; timeout 900 python bench.py --steps 10 --warmup 5 > $OUT/bench_1gpu.json 2> $OUT/bench_1gpu.err; log "exit $?"
  for v in "X=0" "ELB200_F64_WS=1" "ELB200_F64_TMA=1"; do
    log "=== f64 variant $v"
    env $v timeout 120 tools/gemm_check correctness f64 >> $OUT/f64_variants.log 2>&1; log "correctness exit $?"
    env $v timeout 120 tools/native_bench f64 8192 >> $OUT/f64_variants.log 2>&1; log "native_bench exit $?"
    env $v timeout 300 tools/gemm_check perf f64 16384 >> $OUT/f64_variants.log 2>&1; log "perf exit $?"
  done
  # one full ncu capture: the warp-specialised fp64 kernel and cublasDgemm (its launch configuration is the yardstick)
  ELB200_F64_WS=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:'gemm_f64|dgemm|gemm_f64f64|xmma|cutlass' -c 4 \
      -o $OUT/ncu_f64_ws tools/native_bench f64 8192 > $OUT/ncu_f64_ws.log 2>&1; log "ncu f64 exit $?"
  timeout 600 ncu --set full --clock-control none -c 12 -o $OUT/ncu_f64_all tools/native_bench f64 4096 > $OUT/ncu_f64_all.log 2>&1; log "ncu f64 all exit $?"
  ;;
1b)
  # re-run of the measurement half of 1a with small outputs: ncu reports stay in /tmp on the box, only CSV / text exports come back
  log "=== bench.py"; timeout 900 python bench.py --steps 10 --warmup 5 > $OUT/bench_1gpu.json 2> $OUT/bench_1gpu.err; log "exit $?"
  for v in "X=0" "ELB200_F64_WS=1" "ELB200_F64_TMA=1"; do
    log "=== f64 variant $v"
    echo "##### $v" >> $OUT/f64_variants.log
    env $v timeout 120 tools/gemm_check correctness f64 >> $OUT/f64_variants.log 2>&1; log "correctness exit $?"
    env $v timeout 120 tools/native_bench f64 8192 >> $OUT/f64_variants.log 2>&1; log "native_bench exit $?"
    env $v timeout 300 tools/gemm_check perf f64 16384 >> $OUT/f64_variants.log 2>&1; log "perf exit $?"
  done
  ELB200_F64_WS=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:gemm_f64 -c 1 \
      -o /tmp/ncu_f64_ws tools/native_bench f64 8192 > $OUT/ncu_f64_ws.log 2>&1; log "ncu f64 ws exit $?"
  ncu -i /tmp/ncu_f64_ws.ncu-rep --page raw --csv > $OUT/ncu_f64_ws_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f64_ws.ncu-rep --page details > $OUT/ncu_f64_ws_details.txt 2>/dev/null
  ncu -i /tmp/ncu_f64_ws.ncu-rep --page source --csv > $OUT/ncu_f64_ws_source.csv 2>/dev/null
  timeout 600 ncu --set full --clock-control none -c 14 -o /tmp/ncu_f64_all tools/native_bench f64 4096 > $OUT/ncu_f64_all.log 2>&1; log "ncu f64 all exit $?"
  ncu -i /tmp/ncu_f64_all.ncu-rep --page raw --csv > $OUT/ncu_f64_all_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f64_all.ncu-rep --page details > $OUT/ncu_f64_all_details.txt 2>/dev/null
  du -sh $OUT | tee -a $OUT/session.log
  ;;
1c)
  # fp64 tile shapes: multi-CTA-per-SM tiles (64 x 64 x 16 x 3 CTAs, 128 x 64 x 16 x 2 CTAs) against the one-CTA kernels
  for v in "ELB200_F64_TILE=64" "ELB200_F64_TILE=12864" "ELB200_F64_TILE=128" "ELB200_F64_TILE=128 ELB200_F64_WS=1"; do
    log "=== f64 variant $v"
    echo "##### $v" >> $OUT/f64_variants.log
    env $v timeout 120 tools/gemm_check correctness f64 >> $OUT/f64_variants.log 2>&1; log "correctness exit $?"
    env $v timeout 120 tools/native_bench f64 8192 >> $OUT/f64_variants.log 2>&1; log "native_bench exit $?"
    env $v timeout 300 tools/gemm_check perf f64 16384 >> $OUT/f64_variants.log 2>&1; log "perf exit $?"
  done
  echo "##### TMA" >> $OUT/f64_variants.log
  ELB200_F64_TILE=128 ELB200_F64_TMA=1 timeout 60 tools/native_bench f64 4096 >> $OUT/f64_variants.log 2>&1; log "tma exit $?"
  log "=== pytest kernels (default fp64 tile)"; timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q > $OUT/pytest_kernels.log 2>&1; log "exit $?"; tail -2 $OUT/pytest_kernels.log | tee -a $OUT/session.log
  run timeout 120 tools/el_api_check
  run timeout 120 tools/el_comm_check
  for t in 64 12864; do
    ELB200_F64_TILE=$t timeout 600 ncu --set full --import-source on --clock-control none -k regex:tile_kernel -c 1 \
        -o /tmp/ncu_f64_t$t tools/native_bench f64 8192 > $OUT/ncu_f64_t$t.log 2>&1; log "ncu tile $t exit $?"
    ncu -i /tmp/ncu_f64_t$t.ncu-rep --page raw --csv > $OUT/ncu_f64_t${t}_raw.csv 2>/dev/null
    ncu -i /tmp/ncu_f64_t$t.ncu-rep --page details > $OUT/ncu_f64_t${t}_details.txt 2>/dev/null
  done
  du -sh $OUT | tee -a $OUT/session.log
  ;;
1d)
  for v in "ELB200_F64_TILE=64" "ELB200_F64_TILE=644"; do
    log "=== f64 variant $v"
    echo "##### $v" >> $OUT/f64_variants.log
    env $v timeout 120 tools/gemm_check correctness f64 >> $OUT/f64_variants.log 2>&1; log "correctness exit $?"
    env $v timeout 120 tools/native_bench f64 8192 >> $OUT/f64_variants.log 2>&1; log "native_bench exit $?"
    env $v timeout 300 tools/gemm_check perf f64 16384 >> $OUT/f64_variants.log 2>&1; log "perf exit $?"
  done
  for t in 64 644; do
    ELB200_F64_TILE=$t timeout 600 ncu --set full --import-source on --clock-control none -k regex:tile_kernel -c 1 \
        -o /tmp/ncu_f64_t$t tools/native_bench f64 8192 > $OUT/ncu_f64_t$t.log 2>&1; log "ncu tile $t exit $?"
    ncu -i /tmp/ncu_f64_t$t.ncu-rep --page raw --csv > $OUT/ncu_f64_t${t}_raw.csv 2>/dev/null
    ncu -i /tmp/ncu_f64_t$t.ncu-rep --page source --csv > $OUT/ncu_f64_t${t}_source.csv 2>/dev/null
  done
  log "=== bench.py"; timeout 900 python bench.py --steps 10 --warmup 5 > $OUT/bench_1gpu.json 2> $OUT/bench_1gpu.err; log "exit $?"
  du -sh $OUT | tee -a $OUT/session.log
  ;;
1e)
  # full GPU suite (half kernels, communicator exports, facade drivers), half GEMM rate, host-operand switch A/B, launch list + DRAM traffic
  log "=== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; log "exit $?"; tail -4 $OUT/pytest_gpu.log | tee -a $OUT/session.log
  log "=== f16"; timeout 300 tools/native_bench f16 8192 > $OUT/f16.log 2>&1; log "exit $?"; timeout 300 tools/native_bench f16 16384 >> $OUT/f16.log 2>&1; cat $OUT/f16.log | tee -a $OUT/session.log
  for v in "X=0" "ELB200_HOST_BRANGES=4" "ELB200_HOST_TAPER=0"; do
    echo "=== e2e $v" >> $OUT/e2e_variants.log; env $v timeout 200 tools/native_bench e2e 16384 >> $OUT/e2e_variants.log 2>&1; env $v timeout 300 tools/native_bench e2e 32768 >> $OUT/e2e_variants.log 2>&1
  done
  cat $OUT/e2e_variants.log | tee -a $OUT/session.log
  log "=== launch list"; timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $OUT/launches_bench.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-parity --no-sub > $OUT/ncu_launches.log 2>&1; log "exit $?"
  for dt in f32 f64; do
    timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gemm_f -c 3 --csv \
        --log-file $OUT/ncu_traffic_$dt.csv python bench.py --dtype $dt --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --no-sub > $OUT/ncu_traffic_$dt.log 2>&1; log "ncu traffic $dt exit $?"
  done
  du -sh $OUT | tee -a $OUT/session.log
  ;;
1f)
  # the half GEMM under ncu (it runs at 40 % of cublasGemmEx: where does the stage time go?), fp64 with the wider raster group
  log "=== f16"; timeout 300 tools/native_bench f16 8192 > $OUT/f16.log 2>&1; log "exit $?"; cat $OUT/f16.log | tee -a $OUT/session.log
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:gemm_f16 -c 1 -o /tmp/ncu_f16 tools/native_bench f16 8192 > $OUT/ncu_f16.log 2>&1; log "ncu f16 exit $?"
  ncu -i /tmp/ncu_f16.ncu-rep --page raw --csv > $OUT/ncu_f16_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f16.ncu-rep --page details > $OUT/ncu_f16_details.txt 2>/dev/null
  timeout 600 ncu --set full --clock-control none -k regex:cutlass -c 1 -o /tmp/ncu_f16_cublas tools/native_bench f16 8192 > $OUT/ncu_f16_cublas.log 2>&1; log "ncu f16 cublas exit $?"
  ncu -i /tmp/ncu_f16_cublas.ncu-rep --page details > $OUT/ncu_f16_cublas_details.txt 2>/dev/null
  ncu -i /tmp/ncu_f16_cublas.ncu-rep --page raw --csv > $OUT/ncu_f16_cublas_raw.csv 2>/dev/null
  log "=== f64"; timeout 300 tools/native_bench f64 8192 > $OUT/f64.log 2>&1; timeout 300 tools/gemm_check perf f64 16384 >> $OUT/f64.log 2>&1; grep TFLOP $OUT/f64.log | tee -a $OUT/session.log
  du -sh $OUT | tee -a $OUT/session.log
  ;;
f16dbg)
  # the half GEMM without the relay warp; where the stage time goes: 1 = no C stores, 2 = no loads, 4 = no UMMAs
  log "=== test_gpu_half"; timeout 600 python -m pytest tests/test_gpu_half.py -m gpu -x -q > $OUT/pytest_half.log 2>&1; log "exit $?"; tail -2 $OUT/pytest_half.log | tee -a $OUT/session.log
  for d in 0 1 2 4 3 6; do
    echo "##### ELB200_F16_DEBUG=$d" >> $OUT/f16.log
    ELB200_F16_DEBUG=$d timeout 120 tools/native_bench f16 8192 >> $OUT/f16.log 2>&1
  done
  cat $OUT/f16.log | tee -a $OUT/session.log
  ;;
arrive)
  # cross-CTA arrives without the GPU-scope membar: fp32 correctness + rate, half with 6 / 7 stages
  log "=== gemm_check correctness"; timeout 600 tools/gemm_check correctness > $OUT/correctness.log 2>&1; log "exit $?"; tail -2 $OUT/correctness.log | tee -a $OUT/session.log
  log "=== gemm_check perf f32"; timeout 600 tools/gemm_check perf f32 8192 16384 > $OUT/perf_f32.log 2>&1; log "exit $?"; grep TFLOP $OUT/perf_f32.log | tee -a $OUT/session.log
  for st in 6 7; do
    echo "##### ELB200_F16_STAGES=$st" >> $OUT/f16.log
    ELB200_F16_STAGES=$st timeout 120 tools/native_bench f16 8192 >> $OUT/f16.log 2>&1
    ELB200_F16_STAGES=$st timeout 120 tools/native_bench f16 16384 >> $OUT/f16.log 2>&1
  done
  cat $OUT/f16.log | tee -a $OUT/session.log
  log "=== pytest gemm + half"; timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_half.py -m gpu -x -q > $OUT/pytest.log 2>&1; log "exit $?"; tail -2 $OUT/pytest.log | tee -a $OUT/session.log
  ;;
f16mc)
  # the half GEMM with 4-CTA clusters and the B tile multicast: parity, then on / off
  log "=== test_gpu_half"; timeout 600 python -m pytest tests/test_gpu_half.py -m gpu -x -q > $OUT/pytest_half.log 2>&1; log "exit $?"; tail -4 $OUT/pytest_half.log | tee -a $OUT/session.log
  for mc in 1 0; do
    echo "##### ELB200_F16_MULTICAST=$mc" >> $OUT/f16.log
    ELB200_F16_MULTICAST=$mc timeout 120 tools/native_bench f16 8192 >> $OUT/f16.log 2>&1
    ELB200_F16_MULTICAST=$mc timeout 120 tools/native_bench f16 16384 >> $OUT/f16.log 2>&1
  done
  cat $OUT/f16.log | tee -a $OUT/session.log
  ;;
final8)
  # the round's last multi-GPU session, most important first; a step is skipped once DEADLINE seconds have passed
  DEADLINE=${DEADLINE:-300}
  T0=$(date +%s); el() { echo $(( $(date +%s) - T0 )); }
  late() { if [ $(el) -gt $DEADLINE ]; then log "=== SKIP $1 (t=$(el)s)"; return 0; fi; log "=== $1 (t=$(el)s)"; return 1; }
  late "bench.py N=$N" || { timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $? (t=$(el)s)"; }
  late "dist_sweep c4 c5" || { timeout 300 $TR tools/dist_sweep.py --c4 --c5 --out $OUT/dist_sweep_${N}gpu.json > $OUT/dist_sweep.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c6 default" || { timeout 200 $TR tools/dist_sweep.py --c6 --out $OUT/c6_${N}gpu.json > $OUT/c6.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c6 no AB pipeline" || { ELB200_AB_PIPELINE=0 timeout 200 $TR tools/dist_sweep.py --c6 --out $OUT/c6_${N}gpu_noabpipe.json > $OUT/c6_noabpipe.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c4 small, fused collectives off" || { ELB200_FUSED_BYTES=0 timeout 200 $TR tools/dist_sweep.py --c4 --max-bytes 67108864 --out $OUT/dist_sweep_${N}gpu_nofused.json > $OUT/dist_sweep_nofused.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c4 4-8 GB" || { ELB200_ARENA_MB=4096 timeout 300 $TR tools/dist_sweep.py --c4 --min-bytes 4294967296 --max-bytes 8589934592 --out $OUT/dist_sweep_${N}gpu_large.json > $OUT/dist_sweep_large.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c4 large, gather in one piece" || { ELB200_AG_SPLIT_BYTES=0 timeout 200 $TR tools/dist_sweep.py --c4 --min-bytes 268435456 --out $OUT/dist_sweep_${N}gpu_onepiece.json > $OUT/dist_sweep_onepiece.log 2>&1; log "exit $? (t=$(el)s)"; }
  late "c4 large over NCCL" || { ELB200_P2P=0 timeout 200 $TR tools/dist_sweep.py --c4 --min-bytes 268435456 --out $OUT/dist_sweep_${N}gpu_nccl.json > $OUT/dist_sweep_nccl.log 2>&1; log "exit $? (t=$(el)s)"; }
  grep -h "^C[456] f32" $OUT/dist_sweep.log $OUT/c6.log $OUT/c6_noabpipe.log 2>/dev/null | cut -c1-150 | tee -a $OUT/session.log
  ;;
final1)
  # the round's last 1-GPU session: the whole GPU suite with the final library, the bench line, the ncu evidence
  log "=== pytest -m gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; log "exit $?"; tail -3 $OUT/pytest_gpu.log | tee -a $OUT/session.log
  log "=== bench.py"; timeout 600 python bench.py --steps 10 --warmup 5 > $OUT/bench_1gpu.json 2> $OUT/bench_1gpu.err; log "exit $?"
  timeout 300 ncu --set full --import-source on --clock-control none -k regex:gemm_f32 -c 1 -o /tmp/ncu_f32 tools/gemm_check perf f32 8192 > $OUT/ncu_f32.log 2>&1; log "ncu f32 exit $?"
  ncu -i /tmp/ncu_f32.ncu-rep --page raw --csv > $OUT/ncu_f32_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f32.ncu-rep --page details > $OUT/ncu_f32_details.txt 2>/dev/null
  timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gemm_f -c 3 --csv \
      --log-file $OUT/ncu_traffic_f32.csv python bench.py --dtype f32 --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --no-sub > $OUT/ncu_traffic_f32.log 2>&1; log "ncu traffic f32 exit $?"
  log "=== launch list"; timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $OUT/launches_bench.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu --no-parity --no-sub > $OUT/ncu_launches.log 2>&1; log "exit $?"
  timeout 120 tools/native_bench f16 8192 > $OUT/f16.log 2>&1; timeout 120 tools/native_bench f16 16384 >> $OUT/f16.log 2>&1; cat $OUT/f16.log | tee -a $OUT/session.log
  timeout 200 ncu --set full --import-source on --clock-control none -k regex:gemm_f16 -c 1 -o /tmp/ncu_f16 tools/native_bench f16 8192 > $OUT/ncu_f16.log 2>&1; log "ncu f16 exit $?"
  ncu -i /tmp/ncu_f16.ncu-rep --page raw --csv > $OUT/ncu_f16_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f16.ncu-rep --page details > $OUT/ncu_f16_details.txt 2>/dev/null
  du -sh $OUT | tee -a $OUT/session.log
  ;;
ncu32)
  # the default (cta_group::2) fp32 kernel under ncu --set full; gemm_check launches the cta_group::1 form first
  timeout 300 ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k regex:'gemm_f32_3xtf32_kernel<.int.2' -c 1 -o /tmp/ncu_f32 tools/gemm_check perf f32 8192 > $OUT/ncu_f32.log 2>&1; log "ncu f32 exit $?"
  ncu -i /tmp/ncu_f32.ncu-rep --page raw --csv > $OUT/ncu_f32_raw.csv 2>/dev/null
  ncu -i /tmp/ncu_f32.ncu-rep --page details > $OUT/ncu_f32_details.txt 2>/dev/null
  grep -E "Duration|Cluster Size|Compute \(SM\) Throughput" $OUT/ncu_f32_details.txt | tee -a $OUT/session.log
  ;;
benchN)
  log "=== bench.py N=$N"; timeout 400 $TR bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $?"
  ;;
skew)
  log "=== bench.py N=$N (kernel time per rank)"; timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-sub --no-parity --no-cpu --no-e2e > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $?"
  ;;
zfinal)
  # complex instantiations on a B200: the facade check (no torch), then the GPU test file
  log "=== el_api_check"; timeout 60 tools/el_api_check > $OUT/el_api_check.log 2>&1; log "exit $?"; tail -1 $OUT/el_api_check.log | tee -a $OUT/session.log
  log "=== pytest test_gpu_complex"; timeout 170 python -m pytest tests/test_gpu_complex.py -m gpu -x -q > $OUT/pytest_complex.log 2>&1; log "exit $?"; tail -3 $OUT/pytest_complex.log | tee -a $OUT/session.log
  ;;
zpar)
  run timeout 150 $TR tests/dist_worker.py --only check_complex,check_complex_ref_golden
  ;;
par4)
  run timeout 300 $TR tests/dist_worker.py
  ;;
fused)
  # fused small-message AllGather: parity, then the C4 sweep with it on / off; SUMMA-A / B panel pipeline A/B (C6)
  run timeout 900 $TR tests/dist_worker.py
  if [ "$N" = "2" ]; then run timeout 900 $TR tests/dist_worker.py --height 2; fi
  log "=== dist_sweep c4 (fused gather on)"; timeout 900 $TR tools/dist_sweep.py --c4 --out $OUT/dist_sweep_${N}gpu.json > $OUT/dist_sweep.log 2>&1; log "exit $?"
  log "=== dist_sweep c4 small (fused collectives off)"; ELB200_FUSED_BYTES=0 timeout 900 $TR tools/dist_sweep.py --c4 --max-bytes 67108864 --out $OUT/dist_sweep_${N}gpu_nofused.json > $OUT/dist_sweep_nofused.log 2>&1; log "exit $?"
  log "=== dist_sweep c4 small (fused up to 1 GiB)"; ELB200_FUSED_BYTES=1073741824 timeout 900 $TR tools/dist_sweep.py --c4 --max-bytes 268435456 --out $OUT/dist_sweep_${N}gpu_allfused.json > $OUT/dist_sweep_allfused.log 2>&1; log "exit $?"
  log "=== c6 default"; timeout 900 $TR tools/dist_sweep.py --c6 --out $OUT/c6_${N}gpu.json > $OUT/c6.log 2>&1; log "exit $?"
  log "=== c6 no AB pipeline"; ELB200_AB_PIPELINE=0 timeout 900 $TR tools/dist_sweep.py --c6 --out $OUT/c6_${N}gpu_noabpipe.json > $OUT/c6_noabpipe.log 2>&1; log "exit $?"
  grep -h "C6\|1.0 MiB\|  4.0 MiB\|16.0 MiB\| 64.0 MiB\|256.0 MiB\|1024.0 MiB" $OUT/dist_sweep.log $OUT/dist_sweep_nofused.log $OUT/dist_sweep_allfused.log $OUT/c6.log $OUT/c6_noabpipe.log | tee -a $OUT/session.log
  ;;
multi)
  run timeout 900 $TR tests/dist_worker.py
  run timeout 900 $TR tests/dist_worker.py --level3
  if [ "$N" = "2" ]; then run timeout 900 $TR tests/dist_worker.py --height 2; fi
  log "=== el_comm_check under elrun"; timeout 600 tools/elrun $N tools/el_comm_check > $OUT/el_comm_check.log 2>&1; log "exit $?"; tail -3 $OUT/el_comm_check.log | tee -a $OUT/session.log
  log "=== el_gemm_driver under elrun"; ELB_ID_FILE=/tmp/elb_id_$$ RANK=0 timeout 10 true; timeout 600 tools/elrun $N tools/el_comm_check --gridHeight 1 > $OUT/el_comm_check_h1.log 2>&1; log "exit $?"; tail -2 $OUT/el_comm_check_h1.log | tee -a $OUT/session.log
  log "=== bench.py N=$N"; timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 5 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $?"
  ;;
big)
  # 8 GPUs: parity on 2 x 4, the communicator facade, the bench line, C4 / C5 sweeps incl. the 4 and 8 GB points
  run timeout 900 $TR tests/dist_worker.py
  run timeout 900 $TR tests/dist_worker.py --level3
  log "=== el_comm_check under elrun"; timeout 600 tools/elrun $N tools/el_comm_check > $OUT/el_comm_check.log 2>&1; log "exit $?"; tail -2 $OUT/el_comm_check.log | tee -a $OUT/session.log
  log "=== bench.py N=$N"; timeout 1200 $TR bench.py --gpus $N --steps 10 --warmup 5 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $?"
  log "=== dist_sweep c4 c5"; timeout 900 $TR tools/dist_sweep.py --c4 --c5 --out $OUT/dist_sweep_${N}gpu.json > $OUT/dist_sweep.log 2>&1; log "exit $?"
  log "=== dist_sweep c4 large"; ELB200_ARENA_MB=4096 timeout 900 $TR tools/dist_sweep.py --c4 --min-bytes 4294967296 --max-bytes 8589934592 --out $OUT/dist_sweep_${N}gpu_large.json > $OUT/dist_sweep_large.log 2>&1; log "exit $?"
  ;;
switches)
  # 4 GPUs: the opt-in switches that never met hardware — parity with each on, then its A/B
  log "=== bench.py N=$N"; timeout 1200 $TR bench.py --gpus $N --steps 10 --warmup 5 > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; log "exit $?"
  log "=== dist_sweep baseline"; timeout 900 $TR tools/dist_sweep.py --c4 --c5 --out $OUT/dist_sweep_${N}gpu.json > $OUT/dist_sweep.log 2>&1; log "exit $?"
  log "=== AG_SPLIT parity"; ELB200_AG_SPLIT_MB=4 timeout 900 $TR tests/dist_worker.py > $OUT/dist_worker_agsplit.log 2>&1; log "exit $?"
  log "=== AG_SPLIT sweep"; ELB200_AG_SPLIT_MB=4 timeout 900 $TR tools/dist_sweep.py --c4 --min-bytes 16777216 --out $OUT/dist_sweep_${N}gpu_agsplit.json > $OUT/dist_sweep_agsplit.log 2>&1; log "exit $?"
  log "=== AG_SPLIT no-rotate sweep"; ELB200_AG_SPLIT_MB=4 ELB200_AG_ROTATE=0 timeout 900 $TR tools/dist_sweep.py --c4 --min-bytes 16777216 --out $OUT/dist_sweep_${N}gpu_agsplit_norot.json > $OUT/dist_sweep_agsplit_norot.log 2>&1; log "exit $?"
  log "=== AB_PIPELINE parity"; ELB200_AB_PIPELINE=1 timeout 900 $TR tests/dist_worker.py > $OUT/dist_worker_abpipe.log 2>&1; log "exit $?"
  log "=== DOT_PIPELINE off c5"; ELB200_DOT_PIPELINE=0 timeout 900 $TR tools/dist_sweep.py --c5 --out $OUT/dist_sweep_${N}gpu_nodotpipe.json > $OUT/dist_sweep_nodotpipe.log 2>&1; log "exit $?"
  ;;
*)
  echo "unknown stage $STAGE"; exit 2;;
esac
tail -5 $OUT/session.log
