set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -x -q -m gpu -k "adam or training or checkpoint or full_size" > gpurun_out/v_pytest.log 2>&1; tail -3 gpurun_out/v_pytest.log | cut -c1-200
timeout 120 python tools/adam_region2_probe.py 2>&1 | tail -6
timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err; cut -c1-200 gpurun_out/v_bench.json
