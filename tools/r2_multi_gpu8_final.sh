#!/usr/bin/env bash
# final 8-GPU session (charged 8x): gpurun --gpus 8 --timeout 1500 -- bash tools/r2_multi_gpu8_final.sh
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
S=gpurun_out/final_n8_summary.txt
port=29760
run() { local tag=$1 n=$2; shift 2; port=$((port + 1))
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
      --master-addr 127.0.0.1 --master-port "$port" bench.py --gpus $n --steps 20 --warmup 5 \
      > "gpurun_out/final_$tag.log" 2>&1
  echo "$tag rc=$? $(python tools/bench_brief.py gpurun_out/final_$tag.log)" | tee -a $S; }
run n8_default 8 MM25_X=0
run n8_rs_red 8 MM25_TE_RS=0
run n8_mc 8 MM25_MC=1
run n8_mc_epi0 8 MM25_MC=1 MM25_TE_EPI_KB=0
run n4_default 4 MM25_X=0
port=$((port + 1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
    --master-port "$port" tools/profile_dist.py > gpurun_out/final_n8_timeline.txt 2>&1
echo "timeline rc=$?" | tee -a $S
port=$((port + 1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
    --master-port "$port" tests/dist_gpu_worker.py > gpurun_out/final_dist_worker_world8.log 2>&1
echo "dist worker world 8 rc=$?" | tee -a $S
cat $S
