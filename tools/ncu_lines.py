"""Per-source-line instruction / stall-sample shares from an ncu report (dev tool).
usage: python tools/ncu_lines.py <rep> <kernel regex> [launch idx] [min share]"""
import csv, subprocess, sys, io
rep, rx = sys.argv[1], sys.argv[2]
which = sys.argv[3] if len(sys.argv) > 3 else "1"
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.006
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-id", f"::regex:{rx}:{which}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
cur, hdr, out = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = {k: i for i, k in enumerate(r)}
    elif hdr and len(r) > 10 and r[0].isdigit() and r[2] == "-":
        out.append((cur, int(r[0]), r[1], float(r[hdr["Instructions Executed"]] or 0),
                    float(r[hdr["# Samples"]] or 0), float(r[hdr["Thread Instructions Executed"]] or 0)))
ti = sum(o[3] for o in out); ts = sum(o[4] for o in out)
print(f"total warp-inst {ti:.0f} samples {ts:.0f}")
for f, ln, src, ie, sm, tie in out:
    if ie / ti > thr or sm / ts > thr:
        print(f"{f:>14s}:{ln:<4d} inst {ie/ti:6.2%} samp {sm/ts:6.2%} thr/inst {tie/max(ie,1):5.1f} | {src.strip()[:100]}")
