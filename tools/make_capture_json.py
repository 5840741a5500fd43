"""profiles/<name>_capture.json from an ncu report: the profiler-only counters bench.py quotes for its headline kernel
(DRAM traffic, executed warp instructions, issue-slot activity), with the commit and kernel time they were taken at.
usage: python tools/make_capture_json.py report.ncu-rep out.json [kernel name]"""
import csv, io, json, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
kernel = sys.argv[3] if len(sys.argv) > 3 else "render_kernel"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u, v = rows[0], rows[1], rows[2]
col = {k: i for i, k in enumerate(h)}
def val(name, scale=None):
    x = float(v[col[name]].replace(",", ""))
    unit = u[col[name]]
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(unit, 1)
    return x * mult
commit = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
cap = {"commit": commit, "kernel": kernel, "kernel_name_in_report": v[col["Kernel Name"]],
       "kernel_ms": round(val("gpu__time_duration.sum"), 4),
       "dram_bytes": int(val("dram__bytes_read.sum") + val("dram__bytes_write.sum")),
       "inst_executed": int(val("smsp__inst_executed.sum")),
       "issue_active_pct": round(val("smsp__issue_active.avg.pct_of_peak_sustained_active"), 2),
       "threads_per_inst": round(val("smsp__thread_inst_executed_per_inst_executed.ratio"), 2),
       "registers_per_thread": int(val("launch__registers_per_thread")),
       "command": "ncu --set full --clock-control none --import-source on -k regex:render_ -s 1 -c 1 python tools/perf_one.py C"}
json.dump(cap, open(out, "w"), indent=1)
print(json.dumps(cap))
