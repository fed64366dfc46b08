#!/usr/bin/env bash
# 8-GPU session (charged 8x): gpurun --gpus 8 --timeout 1200 -- bash tools/r2_multi_gpu8.sh
#   cfg2 bench at 8 ranks (kept-E default), timeline of one step, cfg5 (AlignCLIP N=65536 d=1024)
#   at 8 ranks, NCCL parity worker at 4 ranks.
set -u
N=8
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
S=gpurun_out/multi_n8_summary.txt
tr() { local port=$1; shift; timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
        --master-addr 127.0.0.1 --master-port "$port" "$@"; }
tr 29731 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n8_cfg2.log 2>&1
echo "bench cfg2 n=8 rc=$?" | tee -a $S
tr 29732 tools/profile_dist.py > gpurun_out/timeline_n8.txt 2>&1
echo "timeline rc=$?" | tee -a $S
tr 29733 bench.py --gpus $N --config cfg5 --steps 10 --warmup 3 > gpurun_out/bench_n8_cfg5.log 2>&1
echo "bench cfg5 n=8 rc=$?" | tee -a $S
N=4
tr 29734 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_n4_cfg2.log 2>&1
echo "bench cfg2 n=4 rc=$?" | tee -a $S
timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x > gpurun_out/dist_tests_n8box.log 2>&1
echo "dist_tests (world 4) rc=$?" | tee -a $S
cat $S
