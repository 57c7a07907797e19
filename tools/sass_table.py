"""Registers / spills (ptxas -v) and SASS instruction count (cuobjdump -sass) of every render kernel instantiation.
    python tools/sass_table.py [out.md]          (rebuilds the library with -Xptxas -v; no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package

load_package()
from cuda_ray_tracer_b200 import build as _build

nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
cmd = [nvcc] + _build.NVCC_FLAGS + ["-Xptxas", "-v", "-shared", "-o", _build.LIB] + \
      [os.path.join(_build.CSRC, s) for s in _build.SOURCES] + ["-ldl"]
log = subprocess.run(cmd, capture_output=True, text=True, check=True).stderr
info = {}
cur = None
for ln in log.splitlines():
    m = re.search(r"Compiling entry function '(\w+)'", ln)
    if m:
        cur = m.group(1)
        info[cur] = {}
        continue
    if cur is None:
        continue
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", ln)
    if m and "stack" not in info[cur]:
        info[cur].update(stack=int(m.group(1)), st=int(m.group(2)), ld=int(m.group(3)))
    m = re.search(r"Used (\d+) registers", ln)
    if m:
        info[cur]["regs"] = int(m.group(1))
        cur = None
sass = subprocess.run(["cuobjdump", "-sass", _build.LIB], capture_output=True, text=True).stdout
count = collections.Counter()
fn = None
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        fn = m.group(1)
        continue
    if fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        count[fn] += 1


def pretty(name):
    d = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    d = re.sub(r"\(rt::SceneView.*", "", d).replace("void rt::", "").replace("(int)", "").replace("(bool)", "")
    return d


rows = ["| kernel (W, CTAs/SM, COUNT, CUBIC, GRID) | registers | stack B | spill st / ld B | SASS instructions |", "|---|---|---|---|---|"]
for k in sorted(info, key=pretty):
    if "render" not in k:
        continue
    v = info[k]
    rows.append("| `{}` | {} | {} | {} / {} | {} |".format(pretty(k), v.get("regs"), v.get("stack"), v.get("st"), v.get("ld"), count.get(k, "?")))
out = "\n".join(rows) + "\n"
print(out)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write("# ptxas -v / cuobjdump -sass per render kernel (nvcc 12.9, sm_100a, flags of build.py)\n\n" + out)
