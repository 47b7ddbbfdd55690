cd $GRAFT_REPO_ROOT
for v in b200 loc512 b200 loc512; do
  echo "== $v"; export P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so
  python - <<'PY'
import torch, bench
dev=torch.device('cuda',0); torch.cuda.set_device(dev)
for n,l,kw in ((256,16,{}),(8192,16,dict(want_bytes=1,max_sets=3)),(2048,16,dict(want_bytes=1,max_sets=12))):
    sh=bench.Loc2dShard(dev,n,l,1,**kw)
    best=min(bench.timed_region(sh,40,10,1)/40 for _ in range(3))
    print('loc2d %dx%d %.2f us'%(n,l,best*1e3))
PY
done
unset P2C_LIB
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests17.log 2>&1; tail -5 gpurun_out/r2_gpu_tests17.log | cut -c1-200
grep -n "^E  .*Error\|^E  .*error" gpurun_out/r2_gpu_tests17.log | cut -c1-230
python -c "import __graft_entry__ as g; g.smoke()"
python tools/prof_module.py bwd 4096 64 10
