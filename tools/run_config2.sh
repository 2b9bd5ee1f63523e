#!/usr/bin/env bash
# BASELINE.json config 2 (shrunk in rows/epochs by default): N clients on N GPUs, full-size DDoSClassifier, seq 128,
# batch 16, FedAvg via the fused NVSwitch kernel.  Usage: tools/run_config2.sh [N=2] [ROWS=6000] [EPOCHS=1] [ROUNDS=1]
set -euo pipefail
N="${1:-2}"; ROWS="${2:-6000}"; EPOCHS="${3:-1}"; ROUNDS="${4:-1}"
OUT="/tmp/config2_n${N}"          # checkpoints are 265 MB each: keep them out of gpurun_out
KEEP="gpurun_out/config2_n${N}"
rm -rf "$OUT" "$KEEP"; mkdir -p "$OUT" "$KEEP"
PKG=detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200
python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29533 \
  -m $PKG.cli launch --num-clients "$((N / ${GPC:-1}))" --gpus-per-client "${GPC:-1}" --transport nvswitch --synthetic-rows "$ROWS" --data-path "$OUT/none.csv" \
  --data-fraction 0.5 --local-epochs "$EPOCHS" --rounds "$ROUNDS" --out-dir "$OUT" --event-log "$OUT/events.jsonl" \
  > "$OUT/stdout.log" 2> "$OUT/stderr.log" || { tail -30 "$OUT/stderr.log"; exit 1; }
grep -E "Average Loss|Test Accuracy|federated aggregation" "$OUT/stdout.log" | head -40
ls "$OUT" | head -30
python - <<PY
import json
ev=[json.loads(l) for l in open("$OUT/events.jsonl")]
for e in ev:
    if e["event"] in ("epoch","federated aggregation","model evaluation","model training"):
        print(e["role"], e["event"], {k:(round(v,4) if isinstance(v,float) else v) for k,v in e.items() if k not in ("ts","role","event")})
PY
# keep only the small artefacts
cp "$OUT"/*.csv "$OUT"/events.jsonl "$OUT"/stdout.log "$KEEP"/ 2>/dev/null || true
rm -f "$KEEP"/synthetic_CICIDS2017.csv
for d in "$OUT"/client*_plots*; do [ -d "$d" ] && cp -r "$d" "$KEEP"/; done
du -sh "$KEEP" | cut -f1
