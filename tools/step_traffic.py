"""Summarise an ncu per-kernel capture of one cfg2 step (see the command in the output header) into
profiles/r02_step_kernels_ncu.txt and profiles/r02_traffic.json (read by bench.py for roofline.traffic /
roofline.step_traffic_ratio, labelled with their source):  python tools/step_traffic.py gpurun_out/r2_step_kernels.csv"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hi]
ix = {n: i for i, n in enumerate(h)}
ker = {}
for r in rows[hi + 1:]:
    if len(r) < len(h):
        continue
    ker.setdefault(int(r[ix["ID"]]), {"name": r[ix["Kernel Name"]]})[r[ix["Metric Name"]]] = (r[ix["Metric Value"]], r[ix["Metric Unit"]])
num = lambda v: float(v[0].replace(",", ""))
gb = lambda v: num(v) * {"byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}[v[1]]
ms = lambda v: num(v) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}[v[1]]
step = []
for k in sorted(ker)[-11:]:  # one step = the last 11 launches
    d = ker[k]
    step.append((d["name"][:60], ms(d["gpu__time_duration.sum"]), gb(d["dram__bytes_read.sum"]), gb(d["dram__bytes_write.sum"]),
                 num(d["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"]), num(d["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]),
                 num(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]), num(d["sm__warps_active.avg.pct_of_peak_sustained_active"])))
tt, tr, tw = sum(s[1] for s in step), sum(s[2] for s in step), sum(s[3] for s in step)
ntt = [s for s in step if "k_ntt" in s[0]]
ntt_bytes = sum(s[2] + s[3] for s in ntt) * 1e9
compulsory = 5.75 * 2 ** 20 * 1024
out = ["# LSB_DUAL_STREAM=0 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fp64_cycles_active...,"
       "gpu__dram_throughput...,smsp__issue_active...,sm__warps_active... --clock-control none",
       "#   -k regex:k_ntt|k_bex|k_ks_inner|k_tensor|k_rescale_lift|k_basis -c 60 python bench.py --steps 1 --warmup 3 --no-cpu --no-api --no-secondary",
       "# round 2 final kernels; one cfg2 step = 1024 x CKKS mult_relin+rescale = the last 11 launches, in launch order",
       "# columns: kernel | time ms (cold-cache, serialised) | dram read GB | dram write GB | fp64 pipe % | dram % | issue % | warps active %"]
for s in step:
    out.append("%-62s | %.3f | %.3f | %.3f | %.1f | %.1f | %.1f | %.1f" % s)
out.append("# sum over the step: %.3f ms under ncu, DRAM %.2f GB read + %.2f GB written = %.2f GB per 1024-op step" % (tt, tr, tw, tr + tw))
out.append("# compulsory traffic (SURVEY 8d): 5.75 MiB per op x 1024 = %.2f GB -> step_traffic_ratio = %.2f" % (compulsory / 1e9, (tr + tw) * 1e9 / compulsory))
out.append("# FP64-pipe-busy time summed over the step: %.2f ms of %.2f ms -- the step's other roof: every kernel does its modular "
           "products on the FP64 pipe, so removing DRAM round trips alone cannot take the step below ~%.1f ms" % (
               sum(s[1] * s[4] / 100 for s in step), tt, sum(s[1] * s[4] / 100 for s in step)))
open(os.path.join(ROOT, "profiles", "r02_step_kernels_ncu.txt"), "w").write("\n".join(out) + "\n")
json.dump({"source": "profiles/r02_step_kernels_ncu.txt (ncu capture of one step, not live)",
           "step_dram_bytes": (tr + tw) * 1e9, "step_compulsory_bytes": compulsory, "step_traffic_ratio": (tr + tw) * 1e9 / compulsory,
           "ntt_launches": len(ntt), "ntt_dram_bytes_per_launch": ntt_bytes / len(ntt),
           "fp64_pipe_busy_ms": sum(s[1] * s[4] / 100 for s in step), "step_ms_under_ncu": tt},
          open(os.path.join(ROOT, "profiles", "r02_traffic.json"), "w"), indent=1)
print("\n".join(out))
