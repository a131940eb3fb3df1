"""Turns the captures of tools/profile_round.sh (gpurun_out/<round>_*) into the text summaries kept under profiles/.
    python tools/profile_digest.py r02"""
import collections, csv, json, os, re, subprocess, sys
R = sys.argv[1] if len(sys.argv) > 1 else "r02"
G, P = "gpurun_out", "profiles"


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return dict(zip(rows[0], rows[2])), dict(zip(rows[0], rows[1]))


def kernel_summary(rep, title, units_per_launch, unit_name, alg_bytes):
    v, u = raw_metrics(rep)
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]
    lines = [f"# {title}", f"# ncu --set full --import-source on --clock-control none (one launch; cold-cache, serialised: shares, not absolutes)",
             f"kernel: {v.get('Kernel Name', '')}"]
    for k in keys:
        if k in v:
            lines.append(f"{k:70s} {v[k]} {u.get(k, '')}")
    for k in sorted(v):
        if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
            try:
                if float(v[k]) >= 0.3:
                    lines.append(f"stall {k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):30s} {float(v[k]):.2f} per issue")
            except ValueError:
                pass
    try:
        inst = float(v["smsp__inst_executed.sum"]); 
        lines.append(f"warp instructions per {unit_name}: {inst / units_per_launch:.1f}")
    except Exception:
        pass
    hot = subprocess.run([sys.executable, "tools/ncu_hot.py", rep, "25"], capture_output=True, text=True).stdout
    lines.append("\n# hottest source lines (samples, instructions, top stall reasons)")
    lines.append(hot)
    return "\n".join(lines) + "\n"


open(f"{P}/{R}_fused_k_process_raw_ncu_full.txt", "w").write(
    kernel_summary(f"{G}/{R}_fused.ncu-rep", "k_process_raw<1,0,1>: fused decode + aggregate, C2 slab of 12.5 M events x 64 B (tools/kbench.py)",
                   12_500_000 / 32, "batch of 32 messages", 76))
open(f"{P}/{R}_k_part_drain_ncu_full.txt", "w").write(
    kernel_summary(f"{G}/{R}_drain.ncu-rep", "k_part_drain<0>: fold of the tuple log, C3 shape (1 M keys, histograms, f64; tools/c3bench.py --slabs 12)",
                   1, "launch", 16))
if os.path.exists(f"{G}/{R}_fold.ncu-rep"):
    open(f"{P}/{R}_k_part_fold_ncu_full.txt", "w").write(
        kernel_summary(f"{G}/{R}_fold.ncu-rep", "k_part_fold<1>: slice-mirror fold of the tuple log, C4 shape on one GPU (10 M keys, 2^24 slots, int32, no histograms; tools/c3bench.py --slabs 12 --int --no-hist --keys 10000000)",
                       1, "launch", 16))
for f in (f"{R}_launch_list_summary.txt", f"{R}_launches.csv"):
    open(f"{P}/{f}", "w").write(open(f"{G}/{f}").read())

# C3 traffic per kernel
rows = list(csv.reader(open(f"{G}/{R}_c3_traffic.csv")))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < len(H):
        continue
    d = dict(zip(H, r)); name = re.sub(r"\(.*", "", d["Kernel Name"]).replace("void ", ""); m = d["Metric Name"]
    val = float(d["Metric Value"].replace(",", "")); un = d["Metric Unit"]
    if "byte" in un.lower():
        val *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(un, 1)
    if m == "gpu__time_duration.sum":
        val *= {"ns": 1e-3, "us": 1, "ms": 1e3}.get(un, 1)
    a = agg.setdefault(name, collections.Counter()); a[m] += val; a["n_" + m] += 1
tot = tt = 0
lines = ["# ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python tools/c3bench.py --slabs 16",
         "# C3 shape: 1 M keys, 64-bucket histograms, f64 values; 200 M events streamed once, windows exported and closed behind the watermark"]
for k, a in agg.items():
    b = a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]; tot += b; tt += a["gpu__time_duration.sum"]
    lines.append(f"{k:28s} n={int(a['n_gpu__time_duration.sum']):3d} time={a['gpu__time_duration.sum'] / 1e3:8.3f} ms  dram read={a['dram__bytes_read.sum'] / 1e6:9.1f} MB  write={a['dram__bytes_write.sum'] / 1e6:9.1f} MB")
per = tot / 200e6
lines.append(f"total: {tt / 1e3:.2f} ms of kernels for 200 M events; DRAM bytes per event = {per:.1f} against 76 algorithmic = {per / 76:.2f}x")
open(f"{P}/{R}_c3_launches_and_traffic.txt", "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
# traffic.json for bench.py
fv, _ = raw_metrics(f"{G}/{R}_fused.ncu-rep")
def to_bytes(x, unit):
    return float(x) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
_, fu = raw_metrics(f"{G}/{R}_fused.ncu-rep")
tr = to_bytes(fv["dram__bytes_read.sum"], fu["dram__bytes_read.sum"]) + to_bytes(fv["dram__bytes_write.sum"], fu["dram__bytes_write.sum"])
json.dump({"k_process_raw_dram_bytes_per_launch": tr, "c3_dram_bytes_per_event": per,
           "source": f"profiles/{R}_fused_k_process_raw_ncu_full.txt, profiles/{R}_c3_launches_and_traffic.txt"}, open(f"{P}/traffic.json", "w"), indent=1)
print("fused traffic per launch", tr)
