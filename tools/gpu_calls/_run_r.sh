set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 100 --warmup 10 > gpurun_out/r_bench_n4.json 2> gpurun_out/r_bench_n4.err; python -c "
import json; d=json.load(open('gpurun_out/r_bench_n4.json')); print(d['value'], d['ms_per_step']); f=d['fedavg']; print({k:f[k] for k in ('two_shot_ms','nvls_ms','nvls_local_cast_ms','nccl_allreduce_baseline_ms')}); print(json.dumps(d.get('federated_round'))[:600]); print(json.dumps(d.get('dp'))[:400])"
