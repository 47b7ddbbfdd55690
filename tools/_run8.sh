cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests8.log 2>&1; tail -12 gpurun_out/r2_gpu_tests8.log | cut -c1-200
grep -n "^E  .*Error\|^E  .*error" gpurun_out/r2_gpu_tests8.log | cut -c1-230
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-sweep > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "n2 exit $?"
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python - <<'PY'
import json
for f in ('n1','n2'):
    for line in open('gpurun_out/r2_bench_%s.json'%f):
        if line.startswith('{'):
            d=json.loads(line); print(f, d['n_gpus'], d['steps'], d['ms_per_step'], d['value'], d.get('exchange'))
PY
