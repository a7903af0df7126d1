# Second 8-GPU lease of round 2: the driver's scaling sequence with bench.py's closing HBM-state pass, then config 5 / config 4 again.
cd $GRAFT_REPO_ROOT
O=gpurun_out
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/scale_r2b_n1.json 2> $O/scale_r2b_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > $O/scale_r2b_n$N.json 2> $O/scale_r2b_n$N.err; echo "n$N rc=$?"
done
# what the next tenant finds: a fresh 1-GPU probe on every GPU
for g in 0 1 2 3 4 5 6 7; do
  CUDA_VISIBLE_DEVICES=$g kind-gpu-sim_b200/bin/gpu-probe --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1 | grep -E "triad_calibrate|kind=triad " | tr '\n' ' ' | sed "s/^/after_scale gpu=$g /"; echo
done > $O/after_scale_r2b.log
cat $O/after_scale_r2b.log | grep -o "after_scale gpu=. \|pace_ns=[-0-9]* ms\|gbs=[0-9.]*" | tr '\n' ' '; echo
# config 5 without any NVLink phase in the probe (so that the two runs start from the same HBM state)
timeout 300 python tools/pod_runner.py pods/probe-8gpu.yaml --beside pods/vllm-cpu-pod.yaml -- --no-sweep --no-p2p --no-sharded > $O/config5_r2.json 2> $O/config5_r2.err; echo "config5 rc=$?"
timeout 300 python tools/pod_runner.py pods/probe-1gpu.yaml --replicas 8 --json -- --iters=10 > $O/pods8_concurrent_r2.json 2> $O/pods8_concurrent_r2.err; echo "pods8 rc=$?"
