cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  timeout 400 $TR --nproc-per-node $n --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --no-sweep --no-e2e > gpurun_out/r2_scale2_n$n.json 2> gpurun_out/r2_scale2_n$n.err; echo "n$n exit $?"
done
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline --no-e2e > gpurun_out/r2_scale2_n1.json 2> gpurun_out/r2_scale2_n1.err
timeout 400 $TR --nproc-per-node 8 --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 --no-sweep --no-e2e --no-exchange > gpurun_out/r2_scale2_n8nox.json 2> gpurun_out/r2_scale2_n8nox.err
python - <<'PY'
import json
for n in ('1','2','4','8','8nox'):
    try:
        for line in open('gpurun_out/r2_scale2_n%s.json'%n):
            if line.startswith('{'):
                d=json.loads(line); print(n, d['ms_per_step'], d['value'], d.get('exchange',{}).get('ranks_arrived'))
    except Exception as e: print(n,'ERR',e)
PY
