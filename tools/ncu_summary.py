"""Reads an .ncu-rep (here, on the CPU box: `ncu -i ... --page raw --csv`) and writes one JSON record per DISTINCT
kernel (the slowest instance of each) with the metrics the roofline discussion uses."""
import csv
import io
import json
import subprocess
import sys

KEEP = {
    "gpu__time_duration.sum": "us",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "launch__registers_per_thread": "regs",
    "launch__shared_mem_per_block_dynamic": "smem_dyn_bytes",
    "launch__waves_per_multiprocessor": "waves",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct",
    "dram__bytes_read.sum": "dram_read_bytes",
    "dram__bytes_write.sum": "dram_write_bytes",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pct_active",
    "sm__inst_executed_pipe_tensor.sum": "tensor_inst",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "smsp__cycles_active.avg": "smsp_cycles_active",
}


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    recs = {}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d.get("Kernel Name", "")
        rec = {"name": name}
        for k, short in KEEP.items():
            if k in d and d[k] != "":
                try:
                    rec[short] = float(d[k].replace(",", ""))
                except ValueError:
                    rec[short] = d[k]
        ui = {h: u for h, u in zip(hdr, units)}
        if "us" in rec and ui.get("gpu__time_duration.sum", "").startswith("ns"):
            rec["us"] /= 1e3
        for k in hdr:  # tensor-pipe metric names differ between ncu versions: keep whatever mentions the tensor pipe
            if "pipe_tensor" in k and "pct" in k and d.get(k) not in ("", None):
                rec.setdefault("tensor_metrics", {})[k] = d[k]
        rec["instances"] = recs.get(name, {}).get("instances", 0) + 1
        if name not in recs or rec.get("us", 0) > recs[name].get("us", 0):
            rec["instances"] = recs.get(name, {}).get("instances", 0) + 1
            recs[name] = rec
        else:
            recs[name]["instances"] += 1
    lst = sorted(recs.values(), key=lambda r: -r.get("us", 0))
    json.dump(lst, open(out, "w"), indent=1)
    for r in lst:
        print(f"{r.get('us', 0):9.1f} us  grid {int(r.get('grid', 0)):6d}  dram {r.get('dram_pct', 0):5.1f}%  "
              f"sm {r.get('sm_pct', 0):5.1f}%  x{r['instances']:3d}  {r['name'][:90]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
