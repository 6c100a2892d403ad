#!/usr/bin/env python
"""Summarise gpurun_out/prof_*.ncu-rep (read on the CPU box with `ncu -i`) into profiles/ncu_summary.md."""
import csv
import glob
import io
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PEAKS = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
KEYS = {
    "gpu__time_duration.sum": "duration_us",
    "dram__bytes_read.sum": "dram_read_MB",
    "dram__bytes_write.sum": "dram_write_MB",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
}


def to_unit(val, unit, want):
    v = float(val.replace(",", ""))
    u = unit.lower()
    if want.endswith("_us"):
        return v / 1e3 if u.startswith("ns") else (v * 1e3 if u.startswith("ms") else v)
    if want.endswith("_MB"):
        return {"byte": v / 1e6, "kbyte": v / 1e3, "mbyte": v, "gbyte": v * 1e3}.get(u, v)
    return v


rows = []
reps = sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "ncuraw_*.csv"))) or sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "prof_*.ncu-rep")))
for rep in reps:
    if rep.endswith(".csv"):          # exported on the GPU box by tools/ncu_all.sh (the reports themselves are too large to bring back)
        out = open(rep).read()
    else:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    if len(r) < 3:
        continue
    hdr, units, vals = r[0], r[1], r[2]
    base = os.path.basename(rep)
    d = {"capture": base[7:-4] if base.endswith(".csv") else base[5:-8], "kernel": vals[hdr.index("Kernel Name")][:70] if "Kernel Name" in hdr else "?"}
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS and v:
            try:
                d[KEYS[h]] = to_unit(v, u, KEYS[h])
            except ValueError:
                pass
    rows.append(d)

with open(os.path.join(ROOT, "profiles", "ncu_summary.md"), "w") as f:
    f.write("# ncu --set full captures (1xB200, `tools/ncu_all.sh`; reports read with `ncu -i ... --page raw --csv`)\n\n")
    f.write(f"Roofline denominators (of measured, MEASURED_PEAKS.json): HBM {PEAKS['hbm_gbs']:.0f} GB/s, bf16 {PEAKS['bf16_tflops']:.0f} TFLOP/s burst.\n")
    f.write("Durations are under the profiler (cold caches, serialised) — use them for shares and stall analysis, not as benchmark values.\n\n")
    f.write("| capture | kernel | dur (us) | grid x block | regs | DRAM R/W (MB) | DRAM % | L2 % | tensor pipe % | warps active % |\n|---|---|---|---|---|---|---|---|---|---|\n")
    for d in rows:
        f.write(f"| {d['capture']} | `{d['kernel']}` | {d.get('duration_us', 0):.1f} | {int(d.get('grid', 0))} x {int(d.get('block', 0))} | "
                f"{int(d.get('regs', 0))} | {d.get('dram_read_MB', 0):.1f} / {d.get('dram_write_MB', 0):.1f} | {d.get('dram_pct', 0):.1f} | "
                f"{d.get('l2_pct', 0):.1f} | {d.get('tensor_pipe_pct', 0):.1f} | {d.get('warps_active_pct', 0):.1f} |\n")
print(open(os.path.join(ROOT, "profiles", "ncu_summary.md")).read())
