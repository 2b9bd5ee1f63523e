set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
DP_FULL=1 timeout 200 $R --master-port 29534 tools/mp_dp_check.py > gpurun_out/l_mp_dp4_full.log 2>&1; tail -5 gpurun_out/l_mp_dp4_full.log | cut -c1-250
timeout 150 $R --master-port 29601 tools/timeline_dp.py > gpurun_out/l_timeline.log 2>&1; head -1 gpurun_out/timeline_dp4.txt; grep -E "barrier|adam|fedavg|Fill|1, 1, 1, 4" gpurun_out/timeline_dp4.txt | cut -c1-110
DCAD_BENCH_CHILD=1 timeout 200 $R --master-port 29611 bench.py --gpus 4 --dp 4 --steps 100 --warmup 10 --no-fedavg > gpurun_out/l_dp4.json 2> gpurun_out/l_dp4.err; python -c "
import json; d=json.load(open('gpurun_out/l_dp4.json')); print('dp4', d['value'], d['ms_per_step'])"
DCAD_BENCH_CHILD=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --dp 2 --steps 100 --warmup 10 --no-fedavg > gpurun_out/l_dp2.json 2> gpurun_out/l_dp2.err; python -c "
import json; d=json.load(open('gpurun_out/l_dp2.json')); print('dp2', d['value'], d['ms_per_step'])"
