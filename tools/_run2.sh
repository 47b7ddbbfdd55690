cd $GRAFT_REPO_ROOT
python tools/diag_bone.py 8 256 16 7 > gpurun_out/r2_diag_bone8.txt 2>&1
cat gpurun_out/r2_diag_bone8.txt
rm -f gpurun_out/r2_grad_table2.jsonl
for v in b200 approx ieee; do
  for case in "64 16 5" "256 16 7"; do
    P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/grad_error_table.py $case >> gpurun_out/r2_grad_table2.jsonl 2>/dev/null
  done
  P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python bench.py --steps 240 --warmup 24 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_$v.json 2>gpurun_out/r2_bench_$v.err
  python -c "
import json;d=json.load(open('gpurun_out/r2_bench_$v.json'))
print('$v', d['ms_per_step'], [(e.get('config','')[:24], e.get('ms_per_step')) for e in d['sweep']])"
done
python -m pytest tests -m gpu -q --maxfail=40 --timeout=900 > gpurun_out/r2_gpu_tests2.log 2>&1; tail -15 gpurun_out/r2_gpu_tests2.log
