#!/usr/bin/env bash
# Multi-GPU A/B of the scatter epilogues of the fused GEMM -> reduce-scatter:
#   gpurun --gpus 2 --timeout 1800 -- bash tools/r2_multi_gpu_ab.sh 2
#   gpurun --gpus 8 --timeout 1800 -- bash tools/r2_multi_gpu_ab.sh 8
set -u
N=${1:-2}
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
echo "=== NCCL parity worker (sharded CLIP / AlignCLIP / retrieval / forward_sampled / CLOOB)"
timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -q -x > "gpurun_out/dist_tests_n$N.log" 2>&1
echo "dist_tests rc=$?" | tee -a gpurun_out/summary_multi.txt
port=29720
for mode in 0 1 2; do
  port=$((port + 1))
  MM25_RS_MODE=$mode timeout 600 python -m torch.distributed.run --nnodes=1 \
      --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$port" \
      bench.py --gpus "$N" --steps 20 --warmup 5 > "gpurun_out/bench_n${N}_rs${mode}.log" 2>&1
  echo "bench n=$N rs_mode=$mode rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_n${N}_rs${mode}.log | tail -1)" \
      | tee -a gpurun_out/summary_multi.txt
done
cat gpurun_out/summary_multi.txt
