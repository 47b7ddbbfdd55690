cd $GRAFT_REPO_ROOT
python tools/widen_bench.py 8192 16 2>&1 | grep -E "pck|renorm|mpjpe"
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-sweep"
$B > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "launch list exit $?"
prof() {  # name, kernel regex, skip, command...
  name=$1; rx=$2; skip=$3; shift 3
  "$@" > gpurun_out/plain_$name.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -o gpurun_out/r02_$name -f "$@" > gpurun_out/ncu_$name.log 2>&1
  echo "$name exit $?"
}
prof tile_256x16 tile_kernel 6 python tools/prof_step.py 256 16 6
prof tile_8192x16 tile_kernel 6 python tools/prof_step.py 8192 16 6
prof tile_4096x64 tile_kernel 6 python tools/prof_step.py 4096 64 6
prof fkfwd_4096x64 fk_tile_fwd 6 python tools/prof_module.py fwd 4096 64 6
prof fkbwd_4096x64 fk_tile_bwd 6 python tools/prof_module.py bwd 4096 64 6
prof fkbwd_256x16 fk_tile_bwd 6 python tools/prof_module.py bwd 256 16 6
prof loc2d_8192x16 loc2d_step 6 python tools/prof_module.py loc2d 8192 16 6
ls -la gpurun_out/*.ncu-rep
