#!/bin/bash
# slow_state_cures.sh -- after a collective has put the GPUs into the slow state, what brings them back?
# Measurement = a FRESH 1-GPU run of the host binary per GPU (calibration verdict + triad GB/s), so no
# long-lived context sits on the GPU between steps.  `gpurun --gpus 2 -- bash tools/slow_state_cures.sh`
cd ${GRAFT_REPO_ROOT:-.}
BIN=kind-gpu-sim_b200/bin/gpu-probe
M() {  # label
  for g in 0 1; do
    out=$(CUDA_VISIBLE_DEVICES=$g $BIN --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1)
    pace=$(echo "$out" | grep -o 'triad_calibrate gpu=0 pace_ns=[-0-9]*' | grep -o '[-0-9]*$')
    gbs=$(echo "$out" | grep '^RESULT kind=triad ' | grep -o 'gbs=[0-9.]*' | head -1)
    csum=$(echo "$out" | grep '^RESULT kind=checksum ' | grep -o 'gbs=[0-9.]*' | head -1)
    echo "MEASURE t=$(date +%s.%N | cut -c1-14) after=\"$1\" gpu=$g pace_ns=$pace triad_$gbs checksum_$csum"
  done
}
T() { python tools/slow_state_child.py k4 4 > /dev/null 2>&1; echo "TRIGGER k4 allreduce 4 s rc=$?"; }
M "fresh lease"
T; M "trigger"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 --no-sweep --no-sharded > /dev/null 2>&1; M "host binary: HBM + P2P copies only"
$BIN --max-gpus=2 --bytes=1g --iters=5 --warmup=1 --no-sweep --no-p2p > /dev/null 2>&1; M "host binary: HBM + sharded checksum (NCCL 16 B + lanes exchange)"
$BIN --max-gpus=2 --bytes=64m --iters=2 --warmup=0 --no-p2p --no-sharded > /dev/null 2>&1; M "host binary: allreduce sweep (NCCL plain/registered/symmetric + K4)"
sleep 20; M "20 s idle"
T; M "trigger again"
$BIN --max-gpus=2 --bytes=64m --iters=2 --warmup=0 --no-p2p --no-sharded > /dev/null 2>&1; M "sweep only"
T; M "trigger #3"
sleep 45; M "45 s idle, no process"
python tools/slow_state_child.py memcpy_peer 3 > /dev/null 2>&1; M "3 s of copy-engine peer copies"
python tools/slow_state_child.py connect > /dev/null 2>&1; M "segments + multicast bound and released, no kernel"
