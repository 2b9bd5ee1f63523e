set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 400 python -m pytest tests/test_multigpu_gpu.py -q -m gpu > gpurun_out/g_pytest_8gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g_pytest_8gpu.log
tail -6 gpurun_out/g_pytest_8gpu.log | cut -c1-300
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/mp_fedavg_check.py > gpurun_out/g_mp_fedavg_8.log 2>&1; tail -11 gpurun_out/g_mp_fedavg_8.log | cut -c1-200
DP_FULL=1 DP_GROUP=4 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 tools/mp_dp_check.py > gpurun_out/g_mp_dp_fed2xdp4.log 2>&1; tail -5 gpurun_out/g_mp_dp_fed2xdp4.log | cut -c1-250
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/g_bench_n8.json 2> gpurun_out/g_bench_n8.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/g_bench_n8.json')); print(d['value'], d['ms_per_step']); print(json.dumps(d.get('fedavg'))[:900]); print(json.dumps(d.get('federated_round'))[:700]); print(json.dumps(d.get('dp'))[:900])"
CONFIG3_TIMEOUT=400 timeout 500 bash tools/run_config3.sh 8 10 225745 auto > gpurun_out/g_config3.log 2>&1; tail -60 gpurun_out/g_config3.log | cut -c1-200
