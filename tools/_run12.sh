cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests12.log 2>&1; tail -6 gpurun_out/r2_gpu_tests12.log | cut -c1-200
grep -n "^E  .*Error\|^E  .*error" gpurun_out/r2_gpu_tests12.log | cut -c1-230
python tools/widen_bench.py 8192 16 2>&1 | tee gpurun_out/r2_widen_8192x16.txt
python tools/quick_bench.py b200 2>/dev/null
python tools/prof_module.py loc2d 8192 16 20
python tools/prof_module.py loc2d 256 16 50
python tools/prof_module.py fwd 4096 64 10
python tools/prof_module.py bwd 4096 64 10
python tools/prof_module.py fwd 256 16 50
python tools/prof_module.py bwd 256 16 50
