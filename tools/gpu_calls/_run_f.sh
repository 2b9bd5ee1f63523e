set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_multigpu_gpu.py tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/f_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest_2gpu.log
tail -30 gpurun_out/f_pytest_2gpu.log | cut -c1-400
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/mp_fedavg_check.py > gpurun_out/f_mp_fedavg.log 2>&1; tail -12 gpurun_out/f_mp_fedavg.log | cut -c1-300
DP_FULL=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 tools/mp_dp_check.py > gpurun_out/f_mp_dp_full.log 2>&1; tail -6 gpurun_out/f_mp_dp_full.log | cut -c1-300
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/f_bench_n2.json 2> gpurun_out/f_bench_n2.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/f_bench_n2.json')); print(d['value'], d['ms_per_step'], d.get('dp'))"
W=/tmp/dcad_quality
export QUALITY_REF_TIMEOUT=480
timeout 300 python tools/quality_parity.py prepare --work $W > gpurun_out/f_quality.log 2>&1
timeout 600 python tools/quality_parity.py reference --work $W >> gpurun_out/f_quality.log 2>&1
timeout 300 python tools/quality_parity.py ours --work $W >> gpurun_out/f_quality.log 2>&1
timeout 100 python tools/quality_parity.py report --work $W --out gpurun_out/config2_quality_parity >> gpurun_out/f_quality.log 2>&1
tail -30 gpurun_out/f_quality.log | cut -c1-200
