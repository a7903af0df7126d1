#!/bin/bash
# slow_state_bisect2.sh -- bench.py --gpus 2 leaves the next process 4 % short; a host-binary run clears it.
# Which part of the two-process run?  `gpurun --gpus 2 -- bash tools/slow_state_bisect2.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
M() {
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs"
  done
}
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}" > /dev/null 2>&1; echo "ran rc=$? : ${*:2}"; }
CURE() { $BIN --max-gpus=2 --bytes=1g --iters=3 --warmup=1 > /dev/null 2>&1; M "cure (host binary, 2 GPUs, all phases)"; }
M "fresh lease"
TR 29611 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20 --combine nccl; M "bench --combine nccl (no exchange object at all)"
CURE
TR 29612 tests/mp_exchange_worker.py; M "mp_exchange_worker (IPC exchange + fused kernels + K4, no bench, no NCCL)"
CURE
GPU_PROBE_NO_MULTICAST=1 TR 29613 tests/mp_exchange_worker.py; M "mp_exchange_worker without multicast"
CURE
GPU_PROBE_CONNECT_TWO_GPUS=1 TR 29614 tests/mp_connect_worker.py; M "connect_ipc + close only (one GPU per rank), no kernel"
CURE
GPU_PROBE_NO_MULTICAST=1 TR 29615 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20; M "bench without multicast"
CURE
TR 29616 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20; M "bench (control: expected slow)"
python tools/slow_state_child.py connect > /dev/null 2>&1; M "single-process connect_local + multicast bind + close (is this the cure?)"
