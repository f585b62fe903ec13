"""Per-source-line hot spots of an ncu report: python tools/ncu_lines.py report.ncu-rep [file-substring] [top]
(reads `ncu --page source --print-source sass,cuda --csv`; compile with -lineinfo)."""
import csv, subprocess, sys, io
rep = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, recs = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1]; hdr = None; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and fname and want in fname and len(hdr) > 10 and len(r) >= 40 and r[0].isdigit():
        g = lambda name: float(r[hdr.index(name)] or 0) if r[hdr.index(name)] not in ("-", "") else 0.0
        try: recs.append((fname.split("/")[-1], int(r[0]), r[1], g("Instructions Executed"), g("# Samples"), g("stall_long_sb"), g("Avg. Threads Executed")))
        except ValueError: pass
ti = sum(x[3] for x in recs) or 1; ts = sum(x[4] for x in recs) or 1
print(f"total instructions {ti:.4g}, samples {ts:.0f}")
key = 3 if len(sys.argv) > 4 and sys.argv[4] == "inst" else 4
recs.sort(key=lambda x: -x[key])
for f, ln, src, ins, smp, lsb, thr in recs[:top]:
    print(f"{f}:{ln:4d} inst {100*ins/ti:5.1f}% smp {100*smp/ts:5.1f}% (long_sb {100*lsb/ts:4.1f}%) thr {thr:4.0f} | {src.strip()[:100]}")
