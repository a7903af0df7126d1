# the driver's scaling sequence on one 4-GPU box (N = 1, 2, 4 back to back, its flags), then what the next tenant finds on every GPU
cd $GRAFT_REPO_ROOT
O=gpurun_out
SECONDS=0; timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/scale_r2e_n1.json 2> $O/scale_r2e_n1.err; echo "n1 rc=$? wall ${SECONDS}s"
for N in 2 4; do
  SECONDS=0; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > $O/scale_r2e_n$N.json 2> $O/scale_r2e_n$N.err; echo "n$N rc=$? wall ${SECONDS}s"
done
for g in 0 1 2 3; do
  CUDA_VISIBLE_DEVICES=$g kind-gpu-sim_b200/bin/gpu-probe --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1 | grep -E "triad_calibrate|kind=triad " | tr '\n' ' ' | grep -o "pace_ns=[-0-9]* ms\|gbs=[0-9.]*" | tr '\n' ' ' | sed "s/^/after_scale gpu=$g /"; echo
done | tee $O/after_scale_r2e.log
