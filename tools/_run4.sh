cd $GRAFT_REPO_ROOT
./tools/exp/mufu_test
export P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_clocks.so
P2C_TILE_CLOCKS=1 python tools/tile_clocks.py 256 16 2>&1 | tail -22
P2C_TILE_CLOCKS=2 python tools/tile_clocks.py 8192 16 2>&1 | tail -22
P2C_TILE_CLOCKS=2 python tools/tile_clocks.py 4096 64 2>&1 | tail -22
unset P2C_LIB
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/r2_bench_d.json 2>gpurun_out/r2_bench_d.err; echo "bench exit $?"; tail -3 gpurun_out/r2_bench_d.err
python -c "
import json;d=json.load(open('gpurun_out/r2_bench_d.json'))
print(d['ms_per_step']); [print(e) for e in d['sweep']]"
