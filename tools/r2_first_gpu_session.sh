#!/usr/bin/env bash
# One-GPU checklist for the additions that were written without GPU time:
#   gpurun --timeout 2400 -- bash tools/r2_first_gpu_session.sh
# Every step runs under its own `timeout`, writes to gpurun_out/, and later steps do not depend
# on earlier ones passing (except the opt-in benches, which need their parity test green).
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
run() { local name=$1; shift; echo "=== $name"; ( timeout "$@" ) > "gpurun_out/$name.log" 2>&1; echo "$name rc=$?" | tee -a gpurun_out/summary.txt; }

run 01_pytest_gpu          1500 python -m pytest tests -m gpu -q -x
run 02_retrieval_only       600 python -m pytest tests/test_zz_retrieval_gpu.py -m gpu -q
run 03_exp_scatter_modes    600 env MM25_TEST_EXPERIMENTAL=1 python -m pytest tests/test_zz_retrieval_gpu.py -m gpu -q -k scatter
run 04_exp_two_cta          600 env MM25_TEST_EXPERIMENTAL=1 python -m pytest tests/test_zz_retrieval_gpu.py -m gpu -q -k two_cta
run 04b_exp_pair_suites    1800 env MM25_TEST_EXPERIMENTAL=1 python -m pytest tests/test_zz_retrieval_gpu.py -m gpu -q -k pair_kernels
run 05_bench_default        900 python bench.py --steps 10 --warmup 3
if grep -q "04_exp_two_cta rc=0" gpurun_out/summary.txt; then
  run 06_bench_two_cta      900 env MM25_GEMM_2CTA=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline
fi
if grep -q "04b_exp_pair_suites rc=0" gpurun_out/summary.txt; then
  run 06b_bench_sweep_2cta  900 env MM25_SWEEP_2CTA=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline
  run 06c_bench_both_2cta   900 env MM25_SWEEP_2CTA=1 MM25_GEMM_2CTA=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline
fi
run 07_bench_paths         1200 python tools/bench_paths.py
# launch list of the default bench (after it exited 0 without ncu)
if grep -q "05_bench_default rc=0" gpurun_out/summary.txt; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 \
      --no-cpu-baseline --no-graph > gpurun_out/08_ncu_launches.log 2>&1
  echo "08_ncu_launches rc=$?" | tee -a gpurun_out/summary.txt
fi
cat gpurun_out/summary.txt
tail -n 3 gpurun_out/05_bench_default.log
