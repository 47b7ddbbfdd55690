cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests7.log 2>&1; tail -8 gpurun_out/r2_gpu_tests7.log | cut -c1-200
grep -n "^E  .*Error" gpurun_out/r2_gpu_tests7.log | cut -c1-230
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "n2 exit $?"
timeout 300 $TR --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-sweep --no-exchange > gpurun_out/r2_bench_n2_nox.json 2> gpurun_out/r2_bench_n2_nox.err; echo "n2 nox exit $?"
timeout 300 $TR --master-port 29513 bench.py --gpus 2 --steps 240 --warmup 24 --no-e2e --no-sweep > gpurun_out/r2_bench_n2_240.json 2> gpurun_out/r2_bench_n2_240.err; echo "n2 240 exit $?"
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err
python - <<'PY'
import json
for f in ('n1','n2','n2_nox','n2_240'):
    try:
        d=json.load(open('gpurun_out/r2_bench_%s.json'%f)); print(f, d['n_gpus'], d['ms_per_step'], d['value'], d.get('exchange'))
        for e in d.get('sweep',[]): print('   ', e.get('config','')[:40], e.get('ms_per_step'), e.get('poses_per_s'), e.get('roofline_frac'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -3 gpurun_out/r2_bench_n2.err
