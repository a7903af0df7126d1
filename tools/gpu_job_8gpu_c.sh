# does the closing pass work at 8 GPUs?  one N=8 run (leaves all 8 GPUs disturbed, then restores them), then what the next tenant finds
cd $GRAFT_REPO_ROOT
O=gpurun_out
SECONDS=0; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus 8 --steps 20 --warmup 5 > $O/scale_r2d_n8.json 2> $O/scale_r2d_n8.err; echo "n8 rc=$? wall ${SECONDS}s"
grep -v "^\*\|OMP_NUM\|^$" $O/scale_r2d_n8.err | tail -3
for g in 0 1 2 3 4 5 6 7; do
  CUDA_VISIBLE_DEVICES=$g kind-gpu-sim_b200/bin/gpu-probe --max-gpus=1 --bytes=4g --iters=5 --warmup=2 2>&1 | grep -E "triad_calibrate|kind=triad " | tr '\n' ' ' | grep -o "pace_ns=[-0-9]* ms\|gbs=[0-9.]*" | tr '\n' ' ' | sed "s/^/after_n8 gpu=$g /"; echo
done | tee $O/after_scale_r2d.log
