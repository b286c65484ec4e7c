"""Share of executed instructions and of stall samples per code region of one kernel:
   python tools/ncu_regions.py rep.ncu-rep 0x2490:0x2d40:primary 0x64e0:0x75c0:dual ...
Regions are SASS offset ranges (tools/sass_loops.py prints the loops); the rest is 'other'."""
import csv, io, subprocess, sys
rep = sys.argv[1]
regions = [(int(a, 16), int(b, 16), n) for a, b, n in [x.split(':') for x in sys.argv[2:]]]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]; ix = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
base, agg = None, {}
for r in rows[2:]:
    if len(r) < len(hdr): continue
    a = r[ix["Address"]]
    try: a = int(a, 16) if a.startswith("0x") else int(a)
    except ValueError: continue
    if base is None: base = a
    off = a - base
    name = "other"
    for lo, hi, n in regions:
        if lo <= off <= hi: name = n
    d = agg.setdefault(name, dict(samples=0, inst=0, thr=0, n=0, **{h: 0 for h in stalls}))
    d["samples"] += int(r[ix["# Samples"]] or 0); d["inst"] += int(r[ix["Instructions Executed"]] or 0)
    d["thr"] += int(r[ix["Thread Instructions Executed"]] or 0); d["n"] += 1
    for h in stalls: d[h] += int(r[ix[h]] or 0)
tot = sum(d["samples"] for d in agg.values()); ti = sum(d["inst"] for d in agg.values())
for k, d in agg.items():
    top = sorted(((d[h], h[6:]) for h in stalls if h != "stall_selected" and d[h]), reverse=True)[:6]
    print(f"{k:9s} static={d['n']:5d} inst={d['inst']/ti*100:5.1f}% samples={d['samples']/tot*100:5.1f}% cycles/issue={d['samples']/max(d['stall_selected'],1):5.2f} "
          f"lanes={d['thr']/max(d['inst'],1):4.1f} | " + " ".join(f"{n}:{v/d['samples']*100:.0f}%" for v, n in top))
