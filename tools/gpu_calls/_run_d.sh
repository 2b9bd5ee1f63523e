set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m pytest tests/test_multigpu_gpu.py tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/d_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest_2gpu.log
tail -30 gpurun_out/d_pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/d_bench_n2.json 2> gpurun_out/d_bench_n2.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/d_bench_n2.json; tail -5 gpurun_out/d_bench_n2.err
W=/tmp/dcad_quality
timeout 600 python tools/quality_parity.py prepare --work $W > gpurun_out/d_quality.log 2>&1
timeout 1500 python tools/quality_parity.py reference --work $W >> gpurun_out/d_quality.log 2>&1
timeout 900 python tools/quality_parity.py ours --work $W >> gpurun_out/d_quality.log 2>&1
timeout 300 python tools/quality_parity.py report --work $W --out gpurun_out/config2_quality_parity >> gpurun_out/d_quality.log 2>&1
tail -40 gpurun_out/d_quality.log
