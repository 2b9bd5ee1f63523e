set -e
PKG=detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200
OUT=/tmp/cli_resume; rm -rf $OUT; mkdir -p $OUT
python -m $PKG.cli client --client-id 1 --transport none --synthetic-rows 3000 --data-path $OUT/none.csv --data-fraction 0.5 --local-epochs 1 --rounds 1 --out-dir $OUT --save-state --no-plots --quiet > $OUT/r1.log 2>&1
ls -la $OUT | grep -E "state|model" 
python -m $PKG.cli client --client-id 1 --transport none --synthetic-rows 3000 --data-path $OUT/none.csv --data-fraction 0.5 --local-epochs 1 --rounds 2 --out-dir $OUT --save-state --resume --no-plots > $OUT/r2.log 2>&1
grep -E "resumed|Average Loss|Test Accuracy" $OUT/r2.log | head
python bench.py --steps 20 --warmup 3 --no-graph 2>/dev/null | cut -c40-140
