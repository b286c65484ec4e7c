"""Per-instruction stall samples from an .ncu-rep source page: python tools/ncu_source.py rep [lo hi]
Prints address, SASS, executed count, samples and the dominant stall columns for addresses in [lo,hi]."""
import csv, io, subprocess, sys
rep = sys.argv[1]
lo = int(sys.argv[2], 16) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3], 16) if len(sys.argv) > 3 else 1 << 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {h: 0 for h in stalls}
nsamp = 0
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        a = int(r[ix["Address"]], 16) if r[ix["Address"]].startswith("0x") else int(r[ix["Address"]])
    except ValueError:
        continue
    base = rows[2][ix["Address"]]
    b = int(base, 16) if base.startswith("0x") else int(base)
    off = a - b
    if off < lo or off > hi:
        continue
    s = {h: int(r[ix[h]] or 0) for h in stalls}
    for h in stalls:
        tot[h] += s[h]
    n = int(r[ix["# Samples"]] or 0)
    nsamp += n
    top = sorted(((v, h) for h, v in s.items() if v), reverse=True)[:3]
    print(f"{off:#06x} {r[ix['Source']][:58]:58s} ex={r[ix['Instructions Executed']]:>10s} thr={r[ix['Avg. Threads Executed']]:>5s} n={n:5d} "
          + " ".join(f"{h[6:]}:{v}" for v, h in top))
print("TOTAL samples", nsamp, {h[6:]: v for h, v in sorted(tot.items(), key=lambda x: -x[1]) if v})
