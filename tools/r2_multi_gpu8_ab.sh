#!/usr/bin/env bash
# 8-GPU A/B session (charged 8x): cfg2 bench under the switches of the sharded step.
#   gpurun --gpus 8 --timeout 1500 -- bash tools/r2_multi_gpu8_ab.sh
set -u
N=8
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
S=gpurun_out/ab_n8_summary.txt
port=29740
run() { local tag=$1; shift; port=$((port + 1))
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
      --master-addr 127.0.0.1 --master-port "$port" bench.py --gpus $N --steps 20 --warmup 5 \
      > "gpurun_out/ab_n8_$tag.log" 2>&1
  echo "$tag rc=$? $(python tools/bench_brief.py gpurun_out/ab_n8_$tag.log)" | tee -a $S; }
run default MM25_X=0
run epi0 MM25_TE_EPI_KB=0
run epi80 MM25_TE_EPI_KB=80
run mc MM25_MC=1
run te0_rs0 MM25_TE=0 MM25_RS_MODE=0
run te0_rs1 MM25_TE=0 MM25_RS_MODE=1
run te0_rs2 MM25_TE=0 MM25_RS_MODE=2
port=$((port + 1))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port "$port" tools/profile_dist.py > gpurun_out/ab_n8_timeline_default.txt 2>&1
echo "timeline rc=$?" | tee -a $S
cat $S
