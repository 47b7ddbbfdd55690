cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x --timeout=300 -k "fk_" > gpurun_out/r2_fk_tests.log 2>&1; echo "fk exit $?"
tail -30 gpurun_out/r2_fk_tests.log | cut -c1-250
rm -f gpurun_out/r2_grad_table3.jsonl
for v in m1 m2 m4; do
  P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/grad_error_table.py 256 16 7 >> gpurun_out/r2_grad_table3.jsonl 2>/dev/null
done
python - <<'PY'
import json
for line in open('gpurun_out/r2_grad_table3.jsonl'):
    d=json.loads(line)
    print(d['lib'], ' '.join(d['grad_per_bone_cuda'][:7]), 'loc2d', d['loc_2d']['cuda'])
PY
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests3.log 2>&1; tail -15 gpurun_out/r2_gpu_tests3.log | cut -c1-200
grep -n "^E  .*Error" gpurun_out/r2_gpu_tests3.log | cut -c1-230
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_c.json 2>gpurun_out/r2_bench_c.err; echo "bench exit $?"
python -c "
import json;d=json.load(open('gpurun_out/r2_bench_c.json'))
print(d['ms_per_step']); [print(e) for e in d['sweep']]"
