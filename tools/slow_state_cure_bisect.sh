#!/bin/bash
# slow_state_cure_bisect.sh -- a full 2-GPU host-binary run brings the fast state back; which phase of it does?
# `gpurun --gpus 2 -- bash tools/slow_state_cure_bisect.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
PORT=29640
LASTPACE=0
M() {
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs"
    [ $g -eq 0 ] && LASTPACE=$pace
  done
}
T() {
  PORT=$((PORT + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20 > /dev/null 2>&1
  M "TRIGGER (bench.py --gpus 2 --no-e2e, sweep to 1 MiB)"
}
TRY() {  # label, command...
  if [ "$LASTPACE" -lt 1060 ]; then T; fi
  "${@:2}" > /dev/null 2>&1
  M "$1"
}
M "fresh lease"
T
TRY "host binary 2 GPUs: HBM only (64 MiB arrays)" $BIN --max-gpus=2 --bytes=64m --iters=3 --warmup=1 --no-sweep --no-p2p --no-sharded
TRY "host binary 2 GPUs: HBM only (1 GiB arrays)" $BIN --max-gpus=2 --bytes=1g --iters=3 --warmup=1 --no-sweep --no-p2p --no-sharded
TRY "host binary 2 GPUs: HBM + sharded" $BIN --max-gpus=2 --bytes=64m --iters=3 --warmup=1 --no-sweep --no-p2p
TRY "host binary 2 GPUs: HBM + P2P copies" $BIN --max-gpus=2 --bytes=64m --iters=3 --warmup=1 --no-sweep --no-sharded
TRY "host binary 2 GPUs: HBM + sweep to 1 MiB" $BIN --max-gpus=2 --bytes=64m --iters=3 --warmup=1 --no-p2p --no-sharded --sweep-max-bytes=1m
TRY "host binary 2 GPUs: HBM + sweep to 1 GiB" $BIN --max-gpus=2 --bytes=64m --iters=3 --warmup=1 --no-p2p --no-sharded
TRY "host binary 2 GPUs: everything (1 GiB arrays)" $BIN --max-gpus=2 --bytes=1g --iters=3 --warmup=1
TRY "python: single-process K4 child 2 s" python tools/slow_state_child.py k4 2
TRY "python: copy-engine peer copies 2 s" python tools/slow_state_child.py memcpy_peer 2
