"""Executed-instruction and stall-sample shares per SOURCE LINE of one profiled kernel.

ncu's CSV source page carries per-SASS-instruction counters but no line numbers; nvdisasm -gi carries the line
table (with inline chains) but no counters.  Both list the kernel's instructions in address order, so they are
joined on the instruction offset.
    python tools/ncu_lines.py report.ncu-rep [--so cuda-ray-tracer_b200/librt_b200.so] [--top 40] [--outer]
--outer attributes an instruction to the outermost frame of its inline chain (the line of the kernel body)."""
import argparse
import collections
import csv
import io
import os
import re
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("rep")
ap.add_argument("--so", default=os.path.join(ROOT, "cuda-ray-tracer_b200", "librt_b200.so"))
ap.add_argument("--top", type=int, default=40)
ap.add_argument("--outer", action="store_true")
a = ap.parse_args()

raw = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
kernel = rows[0][1]
hdr = rows[1]
ia, isrc, iex, ismp = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = [r for r in rows[2:] if len(r) > iex]
base = int(body[0][ia], 16)
counters = {int(r[ia], 16) - base: (int(r[iex] or 0), int(r[ismp] or 0), r[isrc].strip()) for r in body}

# mangled-name fragment to find the section: template args of render_wave_kernel<W, MINB, COUNT>
m = re.search(r"(\w+)<([^>]*)>", kernel)
if m:   # mangled template arguments: (int)N -> LiNE, (bool)B -> LbBE
    frag = m.group(1) + "I" + "".join(("Li%sE" if t == "int" else "Lb%sE") % v for t, v in re.findall(r"\((int|bool)\)(\d+)", m.group(2))) + "E"
else:
    frag = re.sub(r"\W.*", "", kernel.split("(")[0].split("::")[-1])

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(a.so)], cwd=tmp, capture_output=True)
lines_of = {}
for cub in sorted(os.listdir(tmp)):
    text = subprocess.run(["nvdisasm", "-c", "-gi", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    if frag not in text:
        continue
    inside, cur, chain_open = False, ("?", 0), False
    for ln in text.splitlines():
        if ln.startswith(".text."):
            inside = frag in ln
            continue
        if not inside:
            continue
        mm = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if mm:
            # an inline chain is printed innermost first, one "//## File" line per frame
            here = (os.path.basename(mm.group(1)), int(mm.group(2)))
            at = re.search(r'inlined at "([^"]+)", line (\d+)', mm.group(3))
            if not chain_open:
                cur = here
            if a.outer:
                cur = (os.path.basename(at.group(1)), int(at.group(2))) if at else here
            chain_open = True
            continue
        mm = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
        if mm:
            lines_of[int(mm.group(1), 16)] = cur
            chain_open = False
    break

by_line = collections.Counter()
smp_line = collections.Counter()
ops_line = collections.defaultdict(collections.Counter)
tot = stot = 0
for off, (ex, smp, src) in counters.items():
    key = lines_of.get(off, ("?", 0))
    by_line[key] += ex
    smp_line[key] += smp
    t = src.split()
    op = (t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "?")).split(".")[0]
    ops_line[key][op] += ex
    tot += ex
    stot += smp
print("kernel", kernel)
print("total warp instructions", tot, " samples", stot, " mapped offsets", len(lines_of))
for key, c in by_line.most_common(a.top):
    print("  {:22s}:{:<4d} {:6.2f} % exec {:6.2f} % samples  {}".format(
        key[0], key[1], 100.0 * c / tot, 100.0 * smp_line[key] / max(stot, 1),
        " ".join("%s:%.1f" % (k, 100.0 * v / c) for k, v in ops_line[key].most_common(5))))
