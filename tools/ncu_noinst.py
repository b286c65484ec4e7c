"""Where a captured kernel waits for instructions: python tools/ncu_noinst.py rep.ncu-rep [N]
Lists the SASS instructions with the most stall_no_inst samples (with the share of all samples), and the total per stall reason."""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
body = [r for r in rows[2:] if len(r) >= len(hdr)]
tot = sum(int(r[col["# Samples"]] or 0) for r in body)
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
print("total samples", tot)
for h in sorted(reasons, key=lambda h: -sum(int(r[col[h]] or 0) for r in body)):
    print(f"  {h:24s} {100 * sum(int(r[col[h]] or 0) for r in body) / tot:6.2f} %")
base = int(body[0][0], 16)
idx = sorted(range(len(body)), key=lambda i: -int(body[i][col["stall_no_inst"]] or 0))[:N]
for i in idx:
    r = body[i]
    print(f"+{int(r[0], 16) - base:06x} no_inst {100 * int(r[col['stall_no_inst']] or 0) / tot:5.2f} %  all {100 * int(r[col['# Samples']] or 0) / tot:5.2f} %  exec {int(r[col['Instructions Executed']] or 0):>12d}  {r[1].strip()[:70]}")
    if i > 0: print(f"          prev: {body[i-1][1].strip()[:80]}")
