#!/usr/bin/env python
"""Turn an ncu report (brought back in gpurun_out/) into the committed summary under profiles/."""
import csv, json, subprocess, sys
rep = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof_kernels.ncu-rep"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/ncu_kernels_summary.md"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
peaks = json.load(open("MEASURED_PEAKS.json")) if __import__("os").path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}
M = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
     "launch__registers_per_thread", "launch__grid_size", "launch__block_size"]
def val(r, c):
    v, u = r[idx[c]].replace(",", ""), units[idx[c]]
    try:
        v = float(v)
    except ValueError:
        return 0.0
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1e3, "us": 1.0, "ns": 1e-3, "s": 1e6}.get(u, 1.0)
    return v * scale
last = {}
for r in rows[2:]:
    last[r[idx["Kernel Name"]].split("(")[0]] = r          # keep the last (warm) instance of each kernel
lines = ["# ncu --set full, per-kernel summary (B200, one GPU)\n",
         f"Source: `{rep}` captured with `ncu --set full --clock-control none --import-source on` on `scripts/ncu_target.py`",
         "(run plain first, exit 0).  ncu serialises kernels and replays each ~40x with cold caches: read the *shares*",
         f"(DRAM %, tensor-pipe %, achieved GB/s), not the absolute durations.  Measured copy peak {peaks['hbm_gbs']:.0f} GB/s (MEASURED_PEAKS.json).\n",
         "| kernel | duration us | DRAM read MB | DRAM write MB | achieved GB/s | frac of measured copy BW | DRAM % (ncu) | tensor pipe % | warps active % | regs | grid x block |",
         "|---|---|---|---|---|---|---|---|---|---|---|"]
for k, r in last.items():
    dur, rd, wr = val(r, M[0]), val(r, M[1]), val(r, M[2])
    gbs = (rd + wr) / max(dur, 1e-9) / 1e3
    lines.append(f"| `{k.replace('void ', '').replace('b200dp::', '')}` | {dur:.1f} | {rd/1e6:.1f} | {wr/1e6:.1f} | {gbs:.0f} | {gbs/peaks['hbm_gbs']:.2f} | "
                 f"{float(r[idx[M[3]]]):.1f} | {float(r[idx[M[4]]]):.1f} | {float(r[idx[M[5]]]):.1f} | {r[idx[M[6]]]} | {r[idx[M[7]]]} x {r[idx[M[8]]]} |")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
