cd $GRAFT_REPO_ROOT
for v in base rot scan pref b200 base b200; do P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/quick_bench.py $v 2>/dev/null; done
timeout 900 python -m pytest tests -m gpu -q --maxfail=40 --timeout=600 > gpurun_out/r2_gpu_tests9.log 2>&1; tail -6 gpurun_out/r2_gpu_tests9.log | cut -c1-200
grep -n "^E  .*Error\|^E  .*error" gpurun_out/r2_gpu_tests9.log | cut -c1-230
