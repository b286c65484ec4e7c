"""Source lines of one SASS address range of a captured kernel, by executed instructions:
    python tools/ncu_region_lines.py rep.ncu-rep 0x5b00 0x8020 [N]
(the range as offsets from the kernel's first instruction, as cuobjdump -sass prints them)."""
import collections, csv, io, subprocess, sys
rep, lo, hi = sys.argv[1], int(sys.argv[2], 16), int(sys.argv[3], 16)
N = int(sys.argv[4]) if len(sys.argv) > 4 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
cur = line = hdr = None
rows = []
for r in csv.reader(io.StringIO(raw)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No": hdr = r
    elif hdr and len(r) >= len(hdr):
        if r[2] == "-":
            line = (cur, r[0], r[1].strip()[:110])
        elif r[2].startswith("0x"):
            rows.append((int(r[2], 16), line, int(r[hdr.index("Instructions Executed")] or 0), int(r[hdr.index("# Samples")] or 0)))
base = min(a for a, *_ in rows)
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0]
for a, line, ex, sm in rows:
    if lo <= a - base < hi:
        agg[line][0] += ex; agg[line][1] += sm; agg[line][2] += 1; tot[0] += ex; tot[1] += sm
allex = sum(r[2] for r in rows)
print(f"range {lo:#x}..{hi:#x}: {100 * tot[0] / allex:.1f} % of the kernel's executed instructions")
for line, (ex, sm, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:N]:
    print(f"{100 * ex / tot[0]:5.1f} % inst {100 * sm / max(tot[1], 1):5.1f} % samples {n:4d} sass  {line[0]}:{line[1]}  {line[2]}")
