#!/bin/bash
# slow_state_bisect.sh -- which part of `bench.py --gpus 2` leaves the NEXT process 4 % short on the triad?
# Measurement = a fresh 1-GPU run of the host binary per GPU.  `gpurun --gpus 2 -- bash tools/slow_state_bisect.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
M() {
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE t=$(date +%s.%N | cut -c1-14) after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs"
  done
}
TR() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 "${@:2}" > /dev/null 2>&1; echo "ran rc=$? : ${*:2}"; }
M "fresh lease"
TR 29601 bench.py --gpus 2 --steps 20 --warmup 5; M "bench.py --gpus 2 (full)"
M "same, measured again"
sleep 30; M "30 s later"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 > /dev/null 2>&1; M "host binary, 2 GPUs, all phases"
TR 29602 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --sweep-log2max 20; M "bench.py --gpus 2 --no-e2e, sweep only to 1 MiB"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 > /dev/null 2>&1; M "host binary again"
TR 29603 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e; M "bench.py --gpus 2 --no-e2e (full sweep)"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 > /dev/null 2>&1; M "host binary again"
TR 29604 tools/slow_state_child.py nccl 5; M "torch NCCL allreduce 1 GiB for 5 s, two processes"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 > /dev/null 2>&1; M "host binary again"
python bench.py --gpus 1 --steps 5 --warmup 3 > /dev/null 2>&1; M "bench.py --gpus 1 with e2e (16 GiB pinned host)"
