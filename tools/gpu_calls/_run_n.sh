set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/n_bench_n8.json 2> gpurun_out/n_bench_n8.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/n_bench_n8.json')); print(d['value'], d['ms_per_step']); f=d['fedavg']; print({k:f[k] for k in ('two_shot_ms','nvls_ms','nvls_local_cast_ms','nccl_allreduce_baseline_ms','vs_nccl_baseline','fraction_of_link_roofline')}); print(json.dumps(d.get('dp'))[:900])"
DP_FULL=1 DP_GROUP=4 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 tools/mp_dp_check.py > gpurun_out/n_mp_dp_fed2xdp4.log 2>&1; tail -4 gpurun_out/n_mp_dp_fed2xdp4.log | cut -c1-200
