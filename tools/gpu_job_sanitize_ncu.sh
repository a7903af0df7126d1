set -x
cd $GRAFT_REPO_ROOT
python tools/sanitize_driver.py > gpurun_out/sanitize_plain.log 2>&1 || { tail -5 gpurun_out/sanitize_plain.log; exit 1; }
tail -1 gpurun_out/sanitize_plain.log
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_driver.py > gpurun_out/sanitize_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_driver" gpurun_out/sanitize_$tool.log | tail -3
done
python tools/profile_driver.py > gpurun_out/plain_driver_r2.log 2>&1 && tail -2 gpurun_out/plain_driver_r2.log
ncu --set full --clock-control none --import-source on -k regex:triad_tma_paced --launch-skip 27 -c 2 -o gpurun_out/prof_r2_triad -f python tools/profile_driver.py > gpurun_out/ncu_r2_triad.log 2>&1; echo "ncu rc=$?"
python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu > gpurun_out/bench_r2b.json 2>gpurun_out/bench_r2b.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu > gpurun_out/ncu_bench_r2.log 2>&1; echo "launch list rc=$?"
