cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 8 4 2; do
  timeout 400 $TR --nproc-per-node $n --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r2_scale_n$n.json 2> gpurun_out/r2_scale_n$n.err; echo "n$n exit $?"
done
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/r2_scale_n1.json 2> gpurun_out/r2_scale_n1.err
python - <<'PY'
import json
for n in (1,2,4,8):
    try:
        for line in open('gpurun_out/r2_scale_n%d.json'%n):
            if line.startswith('{'):
                d=json.loads(line); print(n, d['ms_per_step'], d['value'], 'e2e', d.get('e2e',{}).get('value'), d.get('exchange'))
                for e in d.get('sweep',[]): print('   ', e.get('config','')[:44], e.get('ms_per_step'), e.get('poses_per_s'), e.get('roofline_frac'), e.get('error'))
    except Exception as e: print(n,'ERR',e)
PY
tail -3 gpurun_out/r2_scale_n8.err
