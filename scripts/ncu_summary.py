"""Key metrics of every kernel in an `ncu --set full` report, as text (and the forward kernel's DRAM traffic as JSON).
    python scripts/ncu_summary.py gpurun_out/prof_tc.ncu-rep profiles/r01_tc_kernels_ncu.txt [profiles/r01_fwd_kernel_ncu.json]"""
import csv, io, json, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "launch__grid_size", "launch__registers_per_thread", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.avg.per_cycle_active",
        "sm__cycles_elapsed.max"]
idx = [(w, hdr.index(w)) for w in want if w in hdr]
lines, fwd = [], []
for r in rows[2:]:
    lines.append("----")
    for w, i in idx:
        lines.append(f"  {w} [{units[i]}] = {r[i]}")
    if "mlp_fwd_tc" in r[hdr.index("Kernel Name")]:
        def val(name):
            i = hdr.index(name); v = float(r[i].replace(",", "")); u = units[i]
            return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}.get(u, 1)
        fwd.append((val("dram__bytes_read.sum") + val("dram__bytes_write.sum"), float(r[hdr.index("gpu__time_duration.sum")].replace(",", ""))))
open(out, "w").write("\n".join(lines) + "\n")
if len(sys.argv) > 3 and fwd:
    big = max(fwd)            # the fine pass is the larger launch
    json.dump({"kernel": "mlp_fwd_tc_kernel (fine pass, 655360 sample rows, training: images saved)",
               "dram_bytes_per_launch": big[0], "duration_us_under_ncu": big[1], "source": rep}, open(sys.argv[3], "w"))
print("\n".join(lines))
