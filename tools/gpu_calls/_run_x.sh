set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_multigpu_gpu.py -x -q -m gpu -k "data_parallel" > gpurun_out/x_pytest_dp.log 2>&1; tail -25 gpurun_out/x_pytest_dp.log | cut -c1-300
timeout 120 python -m pytest tests/test_engine_gpu.py -x -q -m gpu -k "checkpoint or resume" > gpurun_out/x_pytest_ck.log 2>&1; tail -3 gpurun_out/x_pytest_ck.log | cut -c1-300
