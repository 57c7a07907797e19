#!/usr/bin/env python
"""Re-emit the reference's eight bundled scene descriptions as data fixtures under scenes/.

The bundled scenes are the benchmark and parity INPUTS named by BASELINE.json's configs (SURVEY.md §2
row 14); /root/reference does not exist on the GPU box, so the inputs have to travel with the repo.
This script parses each /root/reference/scenes/*.yml with PyYAML and writes the same values in one
canonical layout (fixed key order, flow-style vectors, lights after objects).  Values are emitted
with repr() so every number round-trips exactly.  Run once in the build container:
    python tools/import_scenes.py
"""
import glob
import os
import sys

import yaml

SRC = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/scenes"
DST = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes")

OBJECT_KEYS = ["type", "center", "radius", "origin", "normal", "coefficients", "color", "reflection_ratio"]
LIGHT_KEYS = ["type", "direction", "position", "intensity", "color"]
COEFS = ["x3", "y3", "z3", "x2y", "xy2", "x2z", "xz2", "y2z", "yz2", "xyz",
         "x2", "y2", "z2", "xy", "xz", "yz", "x", "y", "z", "c"]


def num(v):
    return repr(v)


def vec(v):
    return "[" + ", ".join(num(c) for c in v) + "]"


def emit_item(node, keys, out):
    first = True
    for k in keys:
        if k not in node:
            continue
        lead = "  - " if first else "    "
        first = False
        v = node[k]
        if k == "coefficients":
            out.append(lead + "coefficients:")
            for c in COEFS:
                if c in v:
                    out.append("      {}: {}".format(c, num(v[c])))
        elif isinstance(v, list):
            out.append("{}{}: {}".format(lead, k, vec(v)))
        else:
            out.append("{}{}: {}".format(lead, k, v if isinstance(v, str) else num(v)))
    unknown = set(node) - set(keys)
    assert not unknown, unknown


def main():
    os.makedirs(DST, exist_ok=True)
    for path in sorted(glob.glob(os.path.join(SRC, "*.yml"))):
        name = os.path.basename(path)
        with open(path) as f:
            d = yaml.safe_load(f)
        out = ["# data fixture: values of the reference's scenes/{} re-emitted by tools/import_scenes.py".format(name)]
        for k in ("width", "height", "fov", "max_reflections"):
            if k in d:
                out.append("{}: {}".format(k, num(d[k])))
        if "bg_color" in d:
            out.append("bg_color: " + vec(d["bg_color"]))
        out.append("objects:")
        for o in d["objects"]:
            emit_item(o, OBJECT_KEYS, out)
        out.append("light_sources:")
        for l in d["light_sources"]:
            emit_item(l, LIGHT_KEYS, out)
        assert not (set(d) - {"width", "height", "fov", "max_reflections", "bg_color", "objects", "light_sources"})
        with open(os.path.join(DST, name), "w") as f:
            f.write("\n".join(out) + "\n")
        # the re-emitted file must parse back to exactly the same document
        with open(os.path.join(DST, name)) as f:
            assert yaml.safe_load(f) == d, name
        print("wrote", name)


if __name__ == "__main__":
    main()
