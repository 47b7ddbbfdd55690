set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
python -m pytest tests -m gpu -q --maxfail=40 --timeout=900 > gpurun_out/r2_gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_gpu_tests.log
tail -60 gpurun_out/r2_gpu_tests.log
for v in b200 ieee dsum both; do
  for case in "64 16 5" "256 16 7"; do
    P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/grad_error_table.py $case >> gpurun_out/r2_grad_table.jsonl 2>> gpurun_out/r2_grad_table.err
  done
done
cat gpurun_out/r2_grad_table.jsonl | cut -c1-700
tail -5 gpurun_out/r2_grad_table.err
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_a.json 2> gpurun_out/r2_bench_a.err; echo "bench exit $?"
cat gpurun_out/r2_bench_a.json; tail -5 gpurun_out/r2_bench_a.err
