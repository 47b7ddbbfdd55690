cd $GRAFT_REPO_ROOT
for v in base b200 base b200; do P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/quick_bench.py $v 2>/dev/null; done
