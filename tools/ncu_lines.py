"""Per-source-line instruction / stall-sample totals from an ncu report (needs -lineinfo and --import-source on).
usage: python tools/ncu_lines.py report.ncu-rep kernel_regex [units] [invocation]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
norm = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
inv = sys.argv[4] if len(sys.argv) > 4 else "1"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--kernel-id", f"::regex:{rx}:{inv}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname = ""; hdr = None; data = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": print(r[1][:90]); continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] in ("", "Kernel Name"): continue
    try:
        ii = hdr.index("Instructions Executed"); sm = hdr.index("# Samples")
        data.append((fname, int(r[0]), r[1].strip()[:100], int(r[ii]), int(r[sm])))
    except (ValueError, IndexError):
        pass
ti = sum(d[3] for d in data); ts = sum(d[4] for d in data)
print(f"total instr {ti} ({ti/norm:.1f} per unit), samples {ts}")
try:
    for d in sorted(data, key=lambda d: -d[4])[:int(sys.argv[5]) if len(sys.argv) > 5 else 40]:
        print(f"{d[3]/norm:9.1f} inst  {100*d[4]/max(ts,1):5.1f}% smp  {d[0]}:{d[1]:<4d} {d[2]}")
except BrokenPipeError:
    pass
