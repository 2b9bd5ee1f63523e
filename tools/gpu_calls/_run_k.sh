set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
DP_FULL=1 timeout 200 $R --master-port 29534 tools/mp_dp_check.py > gpurun_out/k_mp_dp4_full.log 2>&1; tail -6 gpurun_out/k_mp_dp4_full.log | cut -c1-250
timeout 150 $R --master-port 29601 tools/timeline_dp.py > gpurun_out/k_timeline.log 2>&1; head -1 gpurun_out/timeline_dp4.txt; grep -E "barrier|adam|fedavg|Fill|1, 1, 1, 4" gpurun_out/timeline_dp4.txt | cut -c1-110
timeout 500 $R --master-port 29511 bench.py --gpus 4 --steps 100 --warmup 10 > gpurun_out/k_bench_n4.json 2> gpurun_out/k_bench_n4.err; python -c "
import json; d=json.load(open('gpurun_out/k_bench_n4.json')); print(d['value'], d['ms_per_step']); f=d['fedavg']; print({k:f[k] for k in ('two_shot_ms','nvls_ms','nvls_local_cast_ms','nccl_allreduce_baseline_ms','vs_nccl_baseline','fraction_of_link_roofline')}); print(json.dumps(d.get('dp'))[:500])"
