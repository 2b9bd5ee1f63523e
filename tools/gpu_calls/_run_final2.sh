cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_multigpu_gpu.py -x -q -m gpu > gpurun_out/final2_pytest.log 2>&1; tail -3 gpurun_out/final2_pytest.log | cut -c1-300
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 100 --warmup 5 > gpurun_out/final2_bench.json 2> gpurun_out/final2_bench.err; cut -c1-250 gpurun_out/final2_bench.json; tail -2 gpurun_out/final2_bench.err | cut -c1-300
