set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_x_fullsize_gpu.py -x -q -m gpu > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_pytest.log; tail -4 gpurun_out/t_pytest.log | cut -c1-300
timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err; cut -c1-200 gpurun_out/t_bench.json
DCAD_ADAM_ROW_KERNEL=0 timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/t_bench_flat.json 2> gpurun_out/t_bench_flat.err; cut -c1-200 gpurun_out/t_bench_flat.json
