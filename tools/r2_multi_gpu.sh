#!/usr/bin/env bash
# Multi-GPU session:  gpurun --gpus N --timeout 1800 -- bash tools/r2_multi_gpu.sh N [tests]
#   NCCL parity worker (when "tests" is given), bench at N ranks (kept-E default and MM25_TE=0),
#   per-kernel timeline of one graph-replayed step on rank 0.
set -u
N=${1:-2}
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
S=gpurun_out/multi_n${N}_summary.txt
if [ "${2:-}" = "tests" ]; then
  timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -q -x > "gpurun_out/dist_tests_n$N.log" 2>&1
  echo "dist_tests rc=$?" | tee -a $S
fi
port=29720
for te in 1 0; do
  port=$((port + 1))
  MM25_TE=$te timeout 600 python -m torch.distributed.run --nnodes=1 \
      --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$port" \
      bench.py --gpus "$N" --steps 20 --warmup 5 > "gpurun_out/bench_n${N}_te${te}.log" 2>&1
  echo "bench n=$N te=$te rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_n${N}_te${te}.log | tail -1)" | tee -a $S
done
port=$((port + 1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port "$port" tools/profile_dist.py > "gpurun_out/timeline_n${N}.txt" 2>&1
echo "timeline rc=$?" | tee -a $S
cat $S
