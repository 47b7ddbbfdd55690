cd $GRAFT_REPO_ROOT
for v in b200 ieee m2 approx; do P2C_LIB=$PWD/pedestrians_video_2_carla_b200/libp2c_$v.so python tools/diag_loss.py 256 16 7 2>/dev/null; done
