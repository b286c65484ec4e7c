"""List the loops (backward branches) of one kernel in a cubin with their size and instruction mix:
    python tools/sass_loops.py gpuraytracer_b200/csrc/build/pathtrace.cubin pathTrace_smem
Used to check, without a GPU, how many issue slots one primitive test costs."""
import collections, re, subprocess, sys

cubin, fn = sys.argv[1], sys.argv[2]
sass = subprocess.run(["cuobjdump", "-sass", "-fun", fn, cubin], capture_output=True, text=True).stdout
ins = []
for line in sass.splitlines():
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
print(f"{fn}: {len(ins)} instructions, {len(ins) * 16 / 1024:.1f} KB")
addr2idx = {a: i for i, (a, _) in enumerate(ins)}
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s+)?0x([0-9a-f]+)", t)
    if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr2idx:
        j = addr2idx[int(m.group(1), 16)]
        body = [x for _, x in ins[j:i + 1]]
        ops = collections.Counter(re.sub(r"^@!?U?P\d\s+", "", x).split()[0].split(".")[0] for x in body)
        fp = sum(ops[k] for k in ("FFMA", "FMUL", "FADD"))
        print(f"  loop {ins[j][0]:#06x}..{a:#06x}: {len(body):4d} instr ({len(body) * 16} B)  fp32={fp}  "
              + " ".join(f"{k}:{v}" for k, v in ops.most_common(14)))
