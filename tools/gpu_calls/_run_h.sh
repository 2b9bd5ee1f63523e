set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
R="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 200 $R --master-port 29601 tools/timeline_dp.py > gpurun_out/h_timeline.log 2>&1; tail -3 gpurun_out/h_timeline.log | cut -c1-200
B="bench.py --gpus 4 --dp 4 --steps 100 --warmup 10 --no-fedavg"
for cfg in "1 1" "0 1" "1 0" "0 0"; do set -- $cfg
  DCAD_BENCH_CHILD=1 DCAD_DP_EARLY_ADAM=$1 DCAD_DP_SPARSE_EMB=$2 timeout 200 $R --master-port 2961$1$2 $B > gpurun_out/h_dp4_early$1_sparse$2.json 2> gpurun_out/h_dp4_early$1_sparse$2.err
  python -c "import json; d=json.load(open('gpurun_out/h_dp4_early$1_sparse$2.json')); print('early $1 sparse $2', d['value'], d['ms_per_step'])"
done
timeout 500 $R --master-port 29511 bench.py --gpus 4 --steps 100 --warmup 10 > gpurun_out/h_bench_n4.json 2> gpurun_out/h_bench_n4.err; python -c "
import json; d=json.load(open('gpurun_out/h_bench_n4.json')); print(d['value'], d['ms_per_step']); print(json.dumps(d.get('fedavg'))[:500]); print(json.dumps(d.get('dp'))[:400])"
