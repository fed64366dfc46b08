#!/usr/bin/env bash
# last 8-GPU session of the round at HEAD: cfg2 at 8 and 4 ranks, cfg5 at 8 ranks
set -u
mkdir -p gpurun_out
cd "${GRAFT_REPO_ROOT:-.}"
S=gpurun_out/head_n8_summary.txt
port=29860
run() { local tag=$1 n=$2; shift 2; port=$((port + 1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
      --master-addr 127.0.0.1 --master-port "$port" bench.py --gpus $n "$@" \
      > "gpurun_out/head_$tag.log" 2>&1
  echo "$tag rc=$? $(python tools/bench_brief.py gpurun_out/head_$tag.log)" | tee -a $S; }
run n8_cfg2 8 --steps 20 --warmup 5
run n8_cfg5 8 --config cfg5 --steps 10 --warmup 3
run n4_cfg2 4 --steps 20 --warmup 5
cat $S
