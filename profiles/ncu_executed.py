"""Turns the `ncu --metrics smsp__sass_thread_inst_executed_op_{ffma,fmul,fadd}_pred_on.sum,... -k regex:rasterize --csv`
launch list of ONE steady-state step into profiles/executed.json, which bench.py reads to report `frac_executed`
(executed FFMA*2 + FMUL + FADD thread instructions / kernel time / measured FP32 peak) beside the algorithmic fraction.
usage: python profiles/ncu_executed.py gpurun_out/rNN_executed.csv WORKLOAD > profiles/executed.json"""
import csv
import json
import sys

lines = open(sys.argv[1]).read().splitlines()
hi = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[hi:]))
out = {"workload": sys.argv[2] if len(sys.argv) > 2 else "cfg4_1m_1080p", "source": sys.argv[1].split("/")[-1],
       "note": "thread-level instruction counts of one launch (ncu, predicated-on); mufu = thread instructions on the XU "
               "pipe (ex2, rcp and conversions)"}
names = {"rasterize_kernel": "Rasterize", "rasterize_backward_kernel": "RasterizeBackwards"}
key = {"smsp__sass_thread_inst_executed_op_ffma_pred_on.sum": "ffma", "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum": "fmul",
       "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum": "fadd", "smsp__thread_inst_executed_pipe_xu_pred_on.sum": "mufu",
       "smsp__inst_executed.sum": "warp_inst", "smsp__thread_inst_executed.sum": "thread_inst",
       "gpu__time_duration.sum": "ncu_time_us"}
for r in rows:
    k = r["Kernel Name"].replace("void ", "").split("(")[0].split("<")[0]
    if k not in names or r["Metric Name"] not in key:
        continue
    v = float(r["Metric Value"].replace(",", ""))
    if r["Metric Name"] == "gpu__time_duration.sum":
        v = v / 1e3 if r["Metric Unit"] in ("ns", "nsecond") else v
    else:
        v = int(v)
    out.setdefault(names[k], {})[key[r["Metric Name"]]] = v
print(json.dumps(out, indent=1))
