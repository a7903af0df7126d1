#!/bin/bash
# which closing pass of bench.py leaves the next process fast GPUs?  `gpurun --gpus 2 -- bash tools/slow_state_restore_check.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
PORT=29660
M() {
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs"
  done
}
B() { PORT=$((PORT + 1)); python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20 "$@" > /tmp/b.json 2>/tmp/b.err; echo "bench $* rc=$?"; grep -v "^\*\|OMP_NUM\|^$" /tmp/b.err | tail -2; }
M "fresh lease"
for mode in sequential rank0 sequential rank0; do
  B --no-restore; M "control: --no-restore"
  GPU_PROBE_RESTORE_MODE=$mode B; M "closing pass: $mode"
done
