"""Static SASS instruction count of one kernel by innermost source file and by call site in the kernel body.
    python tools/sass_static.py [mangled-name-fragment]"""
import collections, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "cuda-ray-tracer_b200", "librt_b200.so")
frag = sys.argv[1] if len(sys.argv) > 1 else "render_wave_kernelILi256ELi3ELb0ELb1ELb0EE"
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, capture_output=True)
for cub in sorted(os.listdir(tmp)):
    text = subprocess.run(["nvdisasm", "-c", "-gi", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    if frag not in text:
        continue
    inside = False; cur_inner = ("?", 0); cur_outer = ("?", 0); chain_open = False
    inner = collections.Counter(); outer = collections.Counter(); n = 0
    for ln in text.splitlines():
        if ln.startswith(".text."):
            inside = frag in ln
            continue
        if not inside:
            continue
        mm = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if mm:
            here = (os.path.basename(mm.group(1)), int(mm.group(2)))
            at = re.search(r'inlined at "([^"]+)", line (\d+)', mm.group(3))
            if not chain_open:
                cur_inner = here
            cur_outer = (os.path.basename(at.group(1)), int(at.group(2))) if at else here
            chain_open = True
            continue
        if re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln):
            n += 1; inner[cur_inner[0]] += 1; outer[cur_outer] += 1; chain_open = False
    print(cub, "static instructions", n)
    print("by innermost file:", inner.most_common(10))
    print("by line of the kernel body:", outer.most_common(16))
    break
