"""Summarise an ncu report (raw + source pages) into a small text file for profiles/.

    python scripts/summarize_ncu.py gpurun_out/prof.ncu-rep profiles/name.txt "what was profiled"
"""
import csv
import io
import json
import subprocess
import sys

rep, out, note = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
peaks = json.load(open("MEASURED_PEAKS.json")) if __import__("os").path.exists("MEASURED_PEAKS.json") else {}


def page(name):
    r = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True)
    return list(csv.reader(io.StringIO(r.stdout)))


raw = page("raw")
hdr, unit, vals = raw[0], raw[1], raw[2]
m = dict(zip(hdr, vals))
u = dict(zip(hdr, unit))
keys = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]
lines = [f"# ncu --set full summary: {note}", f"# report: {rep}", ""]
for k in keys:
    if k in m:
        lines.append(f"{k:75s} {m[k]} {u.get(k, '')}")
try:
    t_us = float(m["gpu__time_duration.sum"])
    rd, wr = float(m["dram__bytes_read.sum"]), float(m["dram__bytes_write.sum"])
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(u.get("dram__bytes_read.sum", "byte"), 1)
    tscale = {"us": 1e-6, "ms": 1e-3, "ns": 1e-9}.get(u.get("gpu__time_duration.sum", "us"), 1e-6)
    bw = (rd + wr) * scale / (t_us * tscale) / 1e9
    lines.append(f"{'derived: DRAM traffic / duration':75s} {bw:.0f} GB/s"
                 + (f" = {100 * bw / peaks['hbm_gbs']:.0f}% of measured copy bandwidth ({peaks['hbm_gbs']} GB/s)" if peaks else ""))
except Exception as e:  # noqa: BLE001
    lines.append(f"(derived bandwidth unavailable: {e})")
src = page("source")
if len(src) > 2:
    h = src[1]
    ix = {k: i for i, k in enumerate(h)}
    data = src[2:]

    def f(r, k):
        try:
            return float(r[ix[k]])
        except Exception:  # noqa: BLE001
            return 0.0

    tot = sum(f(r, "# Samples") for r in data) or 1.0
    inst = sum(f(r, "Instructions Executed") for r in data)
    lines += ["", f"warp-level instructions executed: {inst:.0f}; stall samples: {tot:.0f}", "top stall locations (SASS):"]
    stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    for r in sorted(data, key=lambda r: -f(r, "# Samples"))[:15]:
        s = sorted(((k, f(r, k)) for k in stalls if f(r, k) > 0), key=lambda kv: -kv[1])[:2]
        lines.append(f"  {100 * f(r, '# Samples') / tot:5.1f}%  {r[ix['Source']][:80]:80s} {s}")
    agg = {}
    for r in data:
        for k in stalls:
            agg[k] = agg.get(k, 0) + f(r, k)
    lines += ["", "stall reasons (share of samples):"]
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]:
        lines.append(f"  {k:28s} {100 * v / tot:5.1f}%")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
