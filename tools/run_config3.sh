#!/usr/bin/env bash
# BASELINE.json config 3: N (=8) clients on N GPUs, R (=10) FedAvg rounds x 1 local epoch each at the published split
# size (225,745-row CSV, 10 % sample per client -> 847 steps per epoch), aggregation = NVSwitch kernel chosen by
# `--transport auto` (NVLS from 4 clients up), started asynchronously and overlapped with writing the local round's
# files and staging the next round's first epoch.  Usage: tools/run_config3.sh [N=8] [ROUNDS=10] [ROWS=225745] [TRANSPORT=auto]
set -euo pipefail
N="${1:-8}"; ROUNDS="${2:-10}"; ROWS="${3:-225745}"; TRANSPORT="${4:-auto}"
OUT="/tmp/config3_n${N}"          # checkpoints are 265 MB each: keep them out of gpurun_out
KEEP="gpurun_out/config3_n${N}_flow"
rm -rf "$OUT" "$KEEP"; mkdir -p "$OUT" "$KEEP"
PKG=detecting_cyber_attacks_with_distilled_large_language_models_in_distributed_networks_b200
python - <<PY
import sys; sys.path.insert(0, ".")
from $PKG.data.synth import write_synthetic_csv
write_synthetic_csv("$OUT/CICIDS2017.csv", $ROWS, seed=0)
PY
T0=$(date +%s.%N)
timeout "${CONFIG3_TIMEOUT:-600}" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29544 \
  -m $PKG.cli launch --num-clients "$N" --transport "$TRANSPORT" --data-path "$OUT/CICIDS2017.csv" \
  --local-epochs 1 --rounds "$ROUNDS" --out-dir "$OUT" --event-log "$OUT/events.jsonl" --no-plots --round-timeout 60 \
  > "$OUT/stdout.log" 2> "$OUT/stderr.log" || { tail -30 "$OUT/stderr.log"; }
T1=$(date +%s.%N)
python - <<PY
import json, statistics
ev = [json.loads(l) for l in open("$OUT/events.jsonl")]
wall = $T1 - $T0
rounds = {}
for e in ev:
    if e["event"] == "federated aggregation":
        rounds.setdefault(e["round"], []).append(e)
summary = {"clients": $N, "rounds": $ROUNDS, "transport": "$TRANSPORT", "launch_wall_s": wall, "per_round": []}
for r in sorted(rounds):
    es = rounds[r]
    summary["per_round"].append({
        "round": r, "clients_reporting": len(es),
        "device_round_ms_max": max((e.get("device_round_ms") or 0) for e in es),
        "launched_after_s_max": max(e.get("launched_after_s", 0) for e in es),
        "overlapped_host_s_median": statistics.median(e.get("overlapped_host_s", 0) for e in es),
        "phase_total_s_max": max(e.get("total_s", e.get("seconds", 0)) for e in es)})
ep = [e for e in ev if e["event"] == "epoch"]
summary["epoch_seconds_median"] = statistics.median(e["seconds"] for e in ep) if ep else None
summary["epoch_samples_per_s_median_per_client"] = statistics.median(e["samples_per_s"] for e in ep) if ep else None
c1 = [e for e in ev if e["role"] == "CLIENT 1"]
summary["client1_round_wall_s"] = None
tr = [e["ts"] for e in c1 if e["event"] == "model training"]
if len(tr) > 1:
    summary["client1_round_wall_s"] = [round(b - a, 3) for a, b in zip(tr, tr[1:])]
summary["client1_epoch_losses"] = [e["loss"] for e in c1 if e["event"] == "epoch"]
mem = [e for e in ev if e["event"] == "round_members"]
summary["members_events"] = len(mem)
json.dump(summary, open("$KEEP/summary.json", "w"), indent=1)
print(json.dumps(summary, indent=1)[:3000])
PY
cp "$OUT"/*.csv "$OUT"/events.jsonl "$OUT"/stdout.log "$KEEP"/ 2>/dev/null || true
rm -f "$KEEP"/CICIDS2017.csv
tail -5 "$OUT/stderr.log" > "$KEEP/stderr_tail.txt" || true
du -sh "$KEEP" | cut -f1
