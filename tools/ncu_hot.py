"""Hot source lines of an ncu report (captured with --set full --import-source on, code built with -lineinfo).
    python tools/ncu_hot.py report.ncu-rep [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
path = ""; hdr = None; lines = []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        path = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if r[0] == "Function Name" or hdr is None:
        continue
    if r[0].isdigit():
        d = dict(zip(hdr[4:], r[4:]))
        try:
            lines.append((int(d["# Samples"]), int(d["Instructions Executed"]), path, int(r[0]), r[1].strip()[:110], d))
        except (ValueError, KeyError):
            pass
tot_s = sum(l[0] for l in lines) or 1; tot_i = sum(l[1] for l in lines) or 1
print(f"total samples {tot_s}, warp instructions {tot_i}")
stalls = [k for k in (hdr or []) if k.startswith("stall_") and "Not Issued" not in k]
for s, i, pth, ln, src, d in sorted(lines, reverse=True)[:top]:
    ss = sorted(((int(d.get(k, 0) or 0), k[6:]) for k in stalls), reverse=True)[:2]
    print(f"{100 * s / tot_s:5.1f}% smp {100 * i / tot_i:5.1f}% ins  {pth}:{ln:<5d} {src}   [{', '.join(f'{k} {v}' for v, k in ss)}]")
