#!/usr/bin/env bash
# generic one-GPU session: each step under its own timeout, logs into gpurun_out/<tag>_*.log
#   gpurun --timeout 1800 -- bash tools/gpu_session.sh <tag> <step> [<step> ...]
set -u
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
TAG=$1; shift
run() { local name=$1; shift; echo "=== $name"; ( timeout "$@" ) > "gpurun_out/${TAG}_$name.log" 2>&1; echo "$name rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt"; }
for step in "$@"; do
  case $step in
    noalloc)   run noalloc 600 python -m pytest tests/test_zzz_no_alloc.py -m gpu -q ;;
    tc)        run tc 900 python -m pytest tests/test_tc_gpu.py -m gpu -q -x ;;
    pytest)    run pytest 1500 python -m pytest tests -m gpu -q -x ;;
    bench)     run bench 600 python bench.py --steps 10 --warmup 3 ;;
    bench20)   run bench20 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline ;;
    bench_te0) run bench_te0 600 env MM25_TE=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline ;;
    bench_2cta) run bench_2cta 600 env MM25_TE=0 MM25_GEMM_2CTA=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline ;;
    cfgs)      for c in cfg1 cfg3 cfg4 cfg5; do run bench_$c 900 python bench.py --config $c --steps 5 --warmup 3; done
               run bench_cfg3_bf16 600 python bench.py --config cfg3 --precision bf16 --steps 5 --warmup 3 --no-cpu-baseline
               run bench_cfg4_bf16 600 python bench.py --config cfg4 --precision bf16 --steps 5 --warmup 3 --no-cpu-baseline
               run bench_cfg4_vmf 600 python bench.py --config cfg4 --vdist vmf --precision bf16 --steps 5 --warmup 3 --no-cpu-baseline
               run bench_cfg4_pn 600 python bench.py --config cfg4 --vdist projected_normal --precision bf16 --steps 5 --warmup 3 --no-cpu-baseline ;;
    newtests)  run newtests 1200 python -m pytest tests/test_config_sizes_gpu.py tests/test_alignclip_gpu.py -m gpu -q ;;
    fused)     run fused 900 python -m pytest tests/test_fused_rank_gpu.py -m gpu -q ;;
    cfg5)      run bench_cfg5 900 python bench.py --config cfg5 --steps 5 --warmup 3 ;;
    exper)     run exper 1800 env MM25_TEST_EXPERIMENTAL=1 python -m pytest tests/test_zz_retrieval_gpu.py -m gpu -q -k "scatter or two_cta or pair_kernels" ;;
    fp32tc)    run fp32tc 900 python -m pytest tests/test_fp32_tc_gpu.py tests/test_topk_gpu.py tests/test_fused_rank_gpu.py -m gpu -q ;;
    cfg4ref)   run bench_cfg4_ref 900 python bench.py --config cfg4 --steps 5 --warmup 3 --no-cpu-baseline
               run bench_cfg1_ref 600 python bench.py --config cfg1 --steps 10 --warmup 3 --no-cpu-baseline ;;
    cfg3ref)   run bench_cfg3_ref 900 python bench.py --config cfg3 --steps 5 --warmup 3 --no-cpu-baseline ;;
    ncunew)    python tools/profile_new_kernels.py > "gpurun_out/${TAG}_new_plain.log" 2>&1 && \
               timeout 1200 ncu --set full --clock-control none --import-source on \
                 -k regex:"sim_rank_tc|topk_slice|topk_merge|sim_fwd_tc_kernel|sim_grad_tc_kernel|gemm_grouped|sim_mod_tc" -s 16 -c 14 -o "gpurun_out/${TAG}_new_prof" \
                 python tools/profile_new_kernels.py > "gpurun_out/${TAG}_ncu_new.log" 2>&1
               echo "ncunew rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt" ;;
    smoke)     run smoke 600 python __graft_entry__.py smoke ;;
    ncurank)   python tools/profile_new_kernels.py > "gpurun_out/${TAG}_rank_plain.log" 2>&1 && \
               timeout 900 ncu --set full --clock-control none --import-source on \
                 -k regex:"sim_rank_tc|topk_slice|topk_merge" -s 4 -c 4 -o "gpurun_out/${TAG}_rank_prof" \
                 python tools/profile_new_kernels.py > "gpurun_out/${TAG}_ncu_rank.log" 2>&1
               echo "ncurank rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt" ;;
    refarm)    run refarm 900 python bench.py --impl reference --steps 3 --warmup 1 ;;
    diag)      run diag 600 python tools/diag_te.py ;;
    diag0)     run diag0 600 env MM25_TE=0 python tools/diag_te.py ;;
    te)        run te 600 python tools/bench_te.py ;;
    te_skipa)  run te_skipa 600 env MM25_TE_DEBUG_SKIP_A=1 python tools/bench_te.py ;;
    te_stats)  run te_stats 600 env MM25_TE_STATS=1 python tools/bench_te.py ;;
    te_pf)     run te_pf 600 env MM25_TE_PF=12 python tools/bench_te.py ;;
    te_stats_skipa) run te_stats_skipa 600 env MM25_TE_STATS=1 MM25_TE_DEBUG_SKIP_A=1 python tools/bench_te.py ;;
    te2)       run te2 600 env MM25_GEMM_2CTA=1 python tools/bench_te.py ;;
    ncu_te)    python tools/bench_te.py 16384 > "gpurun_out/${TAG}_te_plain.log" 2>&1 && \
               timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_te -s 1 -c 1 \
                 -o "gpurun_out/${TAG}_te_prof" python tools/bench_te.py 16384 > "gpurun_out/${TAG}_ncu_te.log" 2>&1
               echo "ncu_te rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt" ;;
    paths)     run paths 1200 python tools/bench_paths.py ;;
    launches)  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > "gpurun_out/${TAG}_plain.log" 2>&1 && \
               timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
                 --log-file "gpurun_out/${TAG}_launches.csv" python bench.py --steps 2 --warmup 3 \
                 --no-cpu-baseline --no-graph > "gpurun_out/${TAG}_ncu_launches.log" 2>&1
               echo "launches rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt" ;;
    ncufull)   python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > "gpurun_out/${TAG}_plain2.log" 2>&1 && \
               timeout 1200 ncu --set full --clock-control none --import-source on \
                 -k regex:"sim_fwd_tc|gemm_te|sim_grad_tc|gemm_grouped" -s 4 -c 3 -o "gpurun_out/${TAG}_prof" \
                 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-graph > "gpurun_out/${TAG}_ncu_full.log" 2>&1
               echo "ncufull rc=$?" | tee -a "gpurun_out/${TAG}_summary.txt" ;;
    *) echo "unknown step $step" ;;
  esac
done
cat "gpurun_out/${TAG}_summary.txt"
