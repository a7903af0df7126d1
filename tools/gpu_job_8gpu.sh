# One 8-GPU lease, everything the round needs from it (each step independent, own timeout).
cd $GRAFT_REPO_ROOT
O=gpurun_out
snap() { nvidia-smi --query-gpu=index,persistence_mode,temperature.gpu,temperature.memory,clocks.sm,clocks.mem,clocks.gr,clocks.video,power.draw,clocks_event_reasons.active --format=csv > $O/clocks_$1_r2.csv 2>&1; }
snap before
nvidia-smi topo -m > $O/topo_8gpu_r2.txt 2>&1
# 1. the driver's scaling sequence, back to back on one box: N = 1, 2, 4, 8 with its flags
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/scale_r2_n1.json 2> $O/scale_r2_n1.err; echo "n1 rc=$?"
for N in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2950$N bench.py --gpus $N --steps 20 --warmup 5 > $O/scale_r2_n$N.json 2> $O/scale_r2_n$N.err; echo "n$N rc=$?"
done
snap after_scale
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --impl reference --gpus 8 --steps 5 --warmup 1 > $O/scale_r2_ref_n8.json 2> $O/scale_r2_ref_n8.err; echo "ref n8 rc=$?"
# 2. the 8-GPU probe pod's command: HBM on 8 GPUs at once, sharded checksum, allreduce sweep (NCCL plain / registered / symmetric / K4), P2P
timeout 300 kind-gpu-sim_b200/bin/gpu-probe --expect-gpus=8 --measured-hbm-gbs=6532.5 > $O/host_8gpu_r2.log 2> $O/host_8gpu_r2.err; echo "host rc=$?"; tail -1 $O/host_8gpu_r2.log
# 3. which algorithm / protocol NCCL picks per size class (one call per size, INIT + TUNING lines only)
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING,REG NCCL_DEBUG_FILE=/tmp/nccl_debug.%p.log timeout 200 kind-gpu-sim_b200/bin/gpu-probe --expect-gpus=8 --bytes=64m --no-p2p --no-sharded --iters=1 --warmup=0 > $O/host_8gpu_nccldebug_r2.log 2>&1; echo "nccl debug rc=$?"
cat /tmp/nccl_debug.*.log 2>/dev/null | grep -E "NCCL version|NVLS|Algo|algo|proto|Proto|channels|Connected|comm 0x.*rank 0 |Register|register" | sed -E 's/^[^ ]+:[0-9]+:[0-9]+ \[[0-9]+\] //' | awk '!seen[$0]++' | head -c 900000 > $O/nccl_debug_r2.txt; wc -l $O/nccl_debug_r2.txt
# 4. config 5: the 8-GPU probe alone, then beside the stand-in for pods/vllm-cpu-pod.yaml
timeout 300 python tools/pod_runner.py pods/probe-8gpu.yaml --beside pods/vllm-cpu-pod.yaml -- --no-sweep --no-p2p > $O/config5_r2.json 2> $O/config5_r2.err; echo "config5 rc=$?"
# 5. config 4: eight concurrent 1-GPU pods
timeout 300 python tools/pod_runner.py pods/probe-1gpu.yaml --replicas 8 --json -- --iters=10 > $O/pods8_concurrent_r2.json 2> $O/pods8_concurrent_r2.err; echo "pods8 rc=$?"
snap after_all
# 6. the multi-GPU tests that a 1-GPU box skips
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_allreduce.py tests/test_gpu_host_binary.py tests/test_zz_no_alloc.py -m gpu -q -x > $O/pytest_8gpu_r2.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_8gpu_r2.log
