"""Summarise an ncu report here (no GPU): key raw metrics per launch + top stall lines of the first launch.
usage: python scripts/ncu_summary.py report.ncu-rep [kernel-regex]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
rx = sys.argv[2] if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_bytes.sum", "sm__cycles_elapsed.max",
        "sm__inst_executed_pipe_xu.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:]:
    print(r[idx["Kernel Name"]][:60], "|", " ".join("%s=%s%s" % (w.split(".")[0].replace("launch__", "").replace("sm__", ""), r[idx[w]], rows[1][idx[w]]) for w in want if w in idx))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["--kernel-name", "regex:" + rx] if rx else []) +
                     ["--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = []; secs.append(cur); continue
    if cur is not None and r:
        cur.append(r)
if secs:
    s = secs[0]; hdr = s[0]; body = s[1:]
    i_s, i_src, i_ex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
    tot = sum(int(r[i_s]) for r in body)
    print("samples", tot, "sass lines", len(body), "warp-instr", sum(int(r[i_ex]) for r in body))
    stall = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = {h: sum(int(r[hdr.index(h)]) for r in body) for h in stall}
    print(" ".join("%s=%d" % (k[6:], v) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:9]))
    for i, r in sorted(enumerate(body), key=lambda x: -int(x[1][i_s]))[:int(sys.argv[3]) if len(sys.argv) > 3 else 18]:
        print("%6d %6s %8s  %s" % (i, r[i_s], r[i_ex], r[i_src][:100]))
