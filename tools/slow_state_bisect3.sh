#!/bin/bash
# slow_state_bisect3.sh -- ingredients of a two-process run, one at a time.  `gpurun --gpus 2 -- bash tools/slow_state_bisect3.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
PORT=29620
M() {
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs"
  done
}
V() {  # backend ingredients
  PORT=$((PORT + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT tools/slow_state_variant.py "$1" "$2" > /tmp/variant.log 2>&1
  rc=$?; [ $rc -ne 0 ] && tail -3 /tmp/variant.log
  M "two processes: backend=$1 ingredients=$2 (rc=$rc)"
  if echo "$LAST" | grep -q 1082; then :; fi
}
CURE() { $BIN --max-gpus=2 --bytes=1g --iters=3 --warmup=1 > /dev/null 2>&1; M "cure"; }
M "fresh lease"
V gloo kernels
V nccl allreduce
V nccl kernels,allreduce
CURE
V gloo kernels,allreduce
CURE
V nccl torchadd,allreduce
CURE
V gloo torchadd,allreduce
CURE
V gloo allreduce
CURE
# one process, both GPUs, kernels then a collective (the host binary's shape, but torch + NCCL via torch is not possible in one process):
$BIN --max-gpus=2 --bytes=2g --iters=3 --warmup=1 --no-sweep --no-p2p > /dev/null 2>&1; M "host binary: HBM + sharded only"
