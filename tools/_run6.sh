cd $GRAFT_REPO_ROOT
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_fused_step.py tests/test_gpu_module_api.py -m gpu -q --timeout=300 > gpurun_out/r2_tests_2gpu.log 2>&1; tail -12 gpurun_out/r2_tests_2gpu.log | cut -c1-220
grep -n "^E  .*Error" gpurun_out/r2_tests_2gpu.log | cut -c1-230
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 exit $?"; tail -5 gpurun_out/r2_bench_n2.err
cat gpurun_out/r2_bench_n2.json
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-sweep --no-cpu-baseline > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; cat gpurun_out/r2_bench_n1.json | cut -c1-300
