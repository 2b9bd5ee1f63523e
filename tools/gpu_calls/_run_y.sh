set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for sp in 1 0 1 0; do
DCAD_SPLIT_LAST_WGRAD=$sp timeout 200 python bench.py --steps 200 --warmup 20 --no-amp-arm > gpurun_out/y_bench_split$sp.json 2> gpurun_out/y_bench.err; python -c "
import json;d=json.load(open('gpurun_out/y_bench_split$sp.json'));print('split $sp', d['ms_per_step'], d['value'], d['gpu_launches'], d['e2e']['value'])"
done
timeout 300 python -m pytest tests/test_engine_gpu.py tests/test_x_fullsize_gpu.py -x -q -m gpu > gpurun_out/y_pytest.log 2>&1; tail -3 gpurun_out/y_pytest.log | cut -c1-300
