cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_ops.py -m gpu -q --timeout=300 2>&1 | tail -5 | cut -c1-200
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 2 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/r2_scale3_n2.json 2> gpurun_out/r2_scale3_n2.err; echo "n2 exit $?"
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline --no-e2e > gpurun_out/r2_scale3_n1.json 2> gpurun_out/r2_scale3_n1.err
timeout 400 $TR --nproc-per-node 2 --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --no-sweep --no-exchange > gpurun_out/r2_scale3_n2nox.json 2> gpurun_out/r2_scale3_n2nox.err
python - <<'PY'
import json
for n in ('1','2','2nox'):
    try:
        for line in open('gpurun_out/r2_scale3_n%s.json'%n):
            if line.startswith('{'):
                d=json.loads(line); print(n, d['ms_per_step'], d['value'], d.get('exchange'))
                for e in d.get('sweep',[]): print('   ', e.get('config','')[:44], e.get('ms_per_step'), e.get('poses_per_s'), e.get('roofline_frac'), e.get('error'))
    except Exception as e: print(n,'ERR',e)
PY
