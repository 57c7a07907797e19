"""Instruction histogram of one kernel from an .ncu-rep source page: share by opcode class and hottest address ranges.
    python tools/ncu_hot.py report.ncu-rep [bucket]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
bucket = int(sys.argv[2]) if len(sys.argv) > 2 else 64
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ops = collections.Counter(); tot = 0; ranges = collections.Counter(); samples = collections.Counter(); text = {}
def opname(t):
    t = t.strip().split()
    if not t: return "?"
    o = t[1] if t[0].startswith("@") and len(t) > 1 else t[0]
    return o.split(".")[0]
for n, r in enumerate(rows[2:]):
    if len(r) <= iex: continue
    ex = int(r[iex] or 0); tot += ex
    ops[opname(r[isrc])] += ex
    ranges[n // bucket] += ex; samples[n // bucket] += int(r[ismp] or 0)
    text.setdefault(n // bucket, []).append(r[isrc])
print("total warp instructions", tot)
for op, c in ops.most_common(25):
    print("  {:10s} {:6.2f} %".format(op, 100.0 * c / tot))
print("hottest instruction ranges ({} instr each): executed share, stall-sample share".format(bucket))
stot = sum(samples.values()) or 1
for b, c in ranges.most_common(12):
    ops_here = collections.Counter(opname(t) for t in text[b])
    print("  #{:5d} {:5.1f} % exec {:5.1f} % samples  {}".format(b * bucket, 100.0 * c / tot, 100.0 * samples[b] / stot,
          " ".join("%s:%d" % kv for kv in ops_here.most_common(6))))
