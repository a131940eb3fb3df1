"""Summarise an .ncu-rep: headline metrics, stall mix, and the hottest source lines.
usage: python tools/ncu_summary.py <file.ncu-rep> [top_n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2:]
def col(name):
    return [(r[hdr.index(name)], units[hdr.index(name)]) for r in vals] if name in hdr else None
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active" ]
for w in want:
    c = col(w)
    if c: print(f"{w:70s} {[x[0][:60] for x in c]} {c[0][1]}")
for i, h in enumerate(hdr):
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") or h.startswith("smsp__average_warp_latency_issue_stalled") :
        v = [r[i] for r in vals]
        try:
            if float(v[0]) >= 0.3: print(f"{h:90s} {v}")
        except Exception: pass
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None; h2 = None; out = []
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No": h2 = r; continue
    if h2 and len(r) == len(h2) and r[0] != "":
        d = dict(zip(h2, r))
        try: out.append((cur, int(r[0]), r[1].strip()[:100], int(d["# Samples"]), int(d["Instructions Executed"]), int(d.get("L1 Wavefronts Shared Excessive") or 0)))
        except Exception: pass
ts = sum(o[3] for o in out) or 1; ti = sum(o[4] for o in out) or 1
print(f"total samples {ts}  warp-instructions {ti/1e6:.1f} M")
for o in sorted(out, key=lambda o: -o[3])[:top]:
    print(f"{o[0]:22s}:{o[1]:4d} samp {o[3]/ts:5.1%} inst {o[4]/ti:5.1%} excess_wf {o[5]/1e6:6.1f}M | {o[2]}")
if len(sys.argv) > 3:
    print("---- by instruction share")
    for o in sorted(out, key=lambda o: -o[4])[:int(sys.argv[3])]:
        print(f"{o[0]:22s}:{o[1]:4d} inst {o[4]/ti:5.1%} ({o[4]*32/12.5e6:6.1f}/ev) samp {o[3]/ts:5.1%} | {o[2]}")
