"""The C++ host above the C ABI (cuda-ray-tracer_b200/host/): YAML scene loader, camera, headless driver.

CPU part: `rt_headless --dump-scene` prints the parsed scene; its Object / LightSource bytes must equal the Python
mirror's (which test_scene.py pins against the reference's own factories) for the bundled scenes, the synthetic
scene and YAML written in other styles; error messages follow src/scene.cpp's wording including line / column.
GPU part: the driver's frame equals the Python host's, byte for byte."""
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
HOST = os.path.join(ROOT, "cuda-ray-tracer_b200", "host")
EXE = os.path.join(HOST, "rt_headless")
SCENES = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]


@pytest.fixture(scope="module")
def exe(pkg):
    from cuda_ray_tracer_b200 import build
    build.build()
    subprocess.run(["make", "-s", "-C", HOST], check=True)
    return EXE


def dump(exe, path, *extra):
    out = subprocess.run([exe, path, "--dump-scene", *extra], capture_output=True, text=True)
    return out.returncode, out.stdout, out.stderr


def expected_dump(sc):
    lines = ["scene {} {} {:.17g} {:.9g} {:.9g} {:.9g} {} {} {}".format(
        sc.px_width, sc.px_height, sc.vertical_fov, *[float(v) for v in sc.bg_color], sc.max_reflections,
        len(sc.objects), len(sc.lights))]
    for o in sc.objects:
        lines.append("object " + o.tobytes().hex())
    for l in sc.lights:
        b = bytearray(48)
        b[0] = int(l["is_spherical"])
        b[8:32] = l["p"].tobytes()
        b[32:44] = l["light_color"].tobytes()
        lines.append("light " + bytes(b).hex())
    return "\n".join(lines) + "\n"


@pytest.mark.parametrize("name", SCENES)
def test_bundled_scenes_parse_to_the_same_bytes(pkg, exe, scenes_dir, name):
    path = os.path.join(scenes_dir, name + ".yml")
    rc, out, err = dump(exe, path)
    assert rc == 0, err
    assert out == expected_dump(pkg.load_from_file(path))
    rc, out, _ = dump(exe, path, "--width", "3840", "--height", "2160", "--max-reflections", "4")
    assert out == expected_dump(pkg.load_from_file(path).with_overrides(width=3840, height=2160, max_reflections=4))


def test_synthetic_scene_yaml(pkg, exe, tmp_path):
    from cuda_ray_tracer_b200 import synth
    desc = synth.describe(n_spheres=300, n_cubics=9, n_lights=5, width=640, height=360)
    p = tmp_path / "synth.yml"
    p.write_text(synth.to_yaml(desc))
    rc, out, err = dump(exe, str(p))
    assert rc == 0, err
    assert out == expected_dump(pkg.load_from_file(str(p)))


OTHER_STYLES = """---
# comments, flow mappings, quoted scalars, sequences at the parent's indentation, blank lines
width: 320      # trailing comment
height: 200
fov: 45.5
max_reflections: 2
bg_color: [ 0.25 , 0.5, 1 ]

objects:
- type: "polynomial"
  coefficients: {x2: 1, y2: 1.5e0, z2: 1, c: -4, xy: .5}
  color: [0.1, 0.2, 0.3]
  reflection_ratio: 0.25
-
  type: 'sphere'
  center:
    - 1
    - -2.5
    - +3
  radius: 0.75
  color: [1, 1, 1]
- {type: plane, normal: [0, 1, 0], origin: [0, -2, 0], color: [0.5, 0.5, 0.5]}
- type: dingDong
  color: [0.8, 0.8, 0.8]
light_sources:
  - type: spherical
    position: [1, 2, 3]
  - {type: directional, direction: [1, -1, 0.5], intensity: 2.5, color: [1, 0.5, 0.25]}
"""


def test_other_yaml_styles(pkg, exe, tmp_path):
    p = tmp_path / "styles.yml"
    p.write_text(OTHER_STYLES)
    rc, out, err = dump(exe, str(p))
    assert rc == 0, err
    assert out == expected_dump(pkg.load_from_file(str(p)))


BASE = "width: 10\nheight: 10\nfov: 30\nobjects:\n  - type: sphere\n    color: [1, 1, 1]\nlight_sources: []\n"

ERRORS = [
    # (yaml, message as src/scene.cpp words it — marks are 1-based line / column)
    ("height: 10\nfov: 30\nobjects: []\nlight_sources: []\n", "Value 'width' undefined, line: 1 column: 1"),
    ("width: abc\nheight: 10\nfov: 30\nobjects: []\nlight_sources: []\n", "Value 'width' is invalid, line: 1 column: 8"),
    ("width: -3\nheight: 10\nfov: 30\nobjects: []\nlight_sources: []\n", "Value 'width' is invalid, line: 1 column: 8"),
    ("width: 10\nheight: 10\nfov: 30\nlight_sources: []\n", "Value 'objects' undefined, line: 1 column: 1"),
    ("width: 10\nheight: 10\nfov: 30\nobjects: 5\nlight_sources: []\n", "Value 'objects' must be a sequence, line: 4 column: 10"),
    (BASE.replace("type: sphere", "type: torus"), "Unknown surface type: 'torus', line: 5 column: 11"),
    (BASE.replace("    color: [1, 1, 1]\n", ""), "Value 'color' undefined, line: 5 column: 5"),
    (BASE.replace("color: [1, 1, 1]", "color: [1, 1]"), "Value 'color' is invalid, line: 6 column: 12"),
    (BASE.replace("color: [1, 1, 1]", "color: [1, 2, 1]"), "Invalid color: (1, 2, 1)"),
    (BASE.replace("    color: [1, 1, 1]", "    radius: -1\n    color: [1, 1, 1]"), "Negative value for sphere radius: -1"),
    (BASE.replace("    color: [1, 1, 1]", "    reflection_ratio: -0.5\n    color: [1, 1, 1]"),
     "Negative value for object reflection ratio: -0.5"),
    (BASE.replace("type: sphere", "type: polynomial"), "Value 'coefficients' undefined, line: 5 column: 5"),
    (BASE.replace("type: sphere", "type: polynomial\n    coefficients: 3"), "Value 'coefficients' must be a mapping, line: 6 column: 19"),
    (BASE.replace("light_sources: []", "light_sources:\n  - type: laser\n"),
     "Light source type must be 'spherical' or 'directional', line: 8 column: 11"),
    (BASE.replace("light_sources: []", "light_sources:\n  - type: directional\n"), "Value 'direction' undefined, line: 8 column: 5"),
    (BASE.replace("light_sources: []", "light_sources:\n  - type: spherical\n    position: [0, 0, 0]\n    intensity: -2\n"),
     "Negative value for light intensity: -2"),
    ("width: 10\nbg_color: [0, 0, 2]\nheight: 10\nfov: 30\nobjects: []\nlight_sources: []\n", "Invalid color: (0, 0, 2)"),
]


@pytest.mark.parametrize("case", range(len(ERRORS)))
def test_error_messages(exe, tmp_path, case):
    text, message = ERRORS[case]
    p = tmp_path / "bad.yml"
    p.write_text(text)
    rc, out, err = dump(exe, str(p))
    assert rc != 0
    assert err.strip() == message


def test_missing_file_and_usage(exe, tmp_path):
    rc, _, err = dump(exe, str(tmp_path / "nope.yml"))
    assert rc != 0 and err.strip() == "Cannot read the file " + str(tmp_path / "nope.yml")
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode != 0 and "Usage" in out.stderr
    p = tmp_path / "broken.yml"
    p.write_text("width: [1, 2\nheight: 3\n")
    rc, _, err = dump(exe, str(p))
    assert rc != 0 and err.startswith("YAML parser error: ")


def test_defaults(exe, tmp_path):
    """scene.cpp:6-7,167-168: max_reflections 5, white background; sphere defaults centre 0 radius 1 (scene.cpp:101-104)."""
    p = tmp_path / "d.yml"
    p.write_text(BASE)
    rc, out, err = dump(exe, str(p))
    assert rc == 0, err
    head = out.splitlines()[0].split()
    assert head[4:8] == ["1", "1", "1", "5"]
    obj = np.frombuffer(bytes.fromhex(out.splitlines()[1].split()[1])[:160], dtype="<f8")
    assert obj[10] == obj[11] == obj[12] == 1.0 and obj[19] == -1.0 and not obj[16:19].any()


@pytest.mark.gpu
def test_headless_driver_frame_equals_python_host(pkg, exe, scenes_dir, tmp_path):
    from cuda_ray_tracer_b200 import binding, camera
    for name, pos, yaw, pitch in (("dingdong", (0.0, 0.0, 0.0), 90.0, 0.0), ("reflection_test", (3.0, 2.0, -5.0), 80.0, -10.0)):
        raw = tmp_path / (name + ".rgba")
        ppm = tmp_path / (name + ".ppm")
        out = subprocess.run([exe, os.path.join(scenes_dir, name + ".yml"), "--width", "640", "--height", "360",
                              "--max-reflections", "4", "--frames", "3", "--position", "{},{},{}".format(*pos),
                              "--yaw", str(yaw), "--pitch", str(pitch), "--raw", str(raw), "--out", str(ppm),
                              "--option", "accel=1", "--option", "band_taper=65"],   # speed options: no pixel changes
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stderr
        assert out.stdout.startswith("FPS: ") and "last render time:" in out.stdout
        sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(width=640, height=360, max_reflections=4)
        with binding.Renderer(sc) as r:
            r.update(camera.camera_matrix(pos, yaw, pitch))
            ref = r.gather()
        got = np.fromfile(raw, dtype=np.uint8).reshape(360, 640, 4)
        # the two hosts build the camera matrix differently (closed form vs LU inverse): allow the last-bit effect
        diff = np.abs(got.astype(int) - ref.astype(int)).max(axis=2)
        assert (diff > 1).mean() <= 1e-3, (name, int(diff.max()), float((diff > 0).mean()))
        data = open(ppm, "rb").read()
        assert data.startswith(b"P6\n640 360\n255\n")
        px = np.frombuffer(data[len(b"P6\n640 360\n255\n"):], dtype=np.uint8).reshape(360, 640, 3)
        assert np.array_equal(px, got[::-1, :, :3])


def _run_driver(exe, scenes_dir, name, tmp_path, tag, w, h, pos, yaw, pitch, extra=()):
    raw = tmp_path / "{}_{}.rgba".format(name, tag)
    out = subprocess.run([exe, os.path.join(scenes_dir, name + ".yml"), "--width", str(w), "--height", str(h),
                          "--max-reflections", "4", "--frames", "2", "--position", "{},{},{}".format(*pos),
                          "--yaw", str(yaw), "--pitch", str(pitch), "--raw", str(raw), "--dump-camera"] + list(extra),
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    cams = [ln.split()[1:] for ln in out.stdout.splitlines() if ln.startswith("camera ")]
    assert len(cams) == 2 and len(cams[-1]) == 16, out.stdout
    m = np.array([float(v) for v in cams[-1]], dtype=np.float64)
    return np.fromfile(raw, dtype=np.uint8).reshape(h, w, 4), m


@pytest.mark.gpu
def test_headless_driver_against_the_oracle(pkg, oracle, exe, scenes_dir, tmp_path):
    """the C++ driver's frame (ray-tracer.cpp:44-58,215-245 without the window) checked against the
    CPU oracle fed with the very 16 doubles the driver handed to update() (--dump-camera) — +-1 LSB, <= 0.1 % beyond —
    on one GPU and with --gpus 2 / 3 (one process, one context and host thread per GPU, fused gather over peer access).
    On a box with a single GPU the ranks share device 0 (option same_device): same code path, same hand-shake."""
    import torch
    from parity import assert_parity
    n_dev = torch.cuda.device_count()
    w, h = 480, 270
    for name, pos, yaw, pitch in (("dingdong", (0.0, 0.0, 0.0), 90.0, 0.0), ("reflection_test", (3.0, 2.0, -5.0), 80.0, -10.0),
                                  ("20spheres", (1.0, 0.5, -2.0), 95.0, 5.0)):
        sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(width=w, height=h, max_reflections=4)
        one, m = _run_driver(exe, scenes_dir, name, tmp_path, "g1", w, h, pos, yaw, pitch)
        want = oracle.quantize_rgba8(oracle.PortOracle(sc).render(m, threads=os.cpu_count()))
        res = assert_parity(one, want, name + " C++ driver, 1 GPU")
        assert res["exact_frac"] > 0.999, res
        if oracle.have_ref():
            rgb, _ = oracle.RefCpu(sc).render_rows(m, 100, 140, 1, threads=1)
            assert_parity(one[100:140], oracle.quantize_rgba8(rgb), name + " C++ driver rows 100..139 (reference build)")
        for gpus in (2, 3):
            extra = ["--gpus", str(gpus), "--option", "strip_rows=8"]
            if n_dev < gpus:
                extra += ["--option", "same_device=1"]
            many, m2 = _run_driver(exe, scenes_dir, name, tmp_path, "g{}".format(gpus), w, h, pos, yaw, pitch, extra)
            assert np.array_equal(m2, m)
            assert np.array_equal(many, one), "{}: --gpus {} frame differs from the 1-GPU frame".format(name, gpus)


def test_headless_driver_rejects_malformed_option(exe, scenes_dir):
    """--option wants name=value; anything else prints the usage text (needs no GPU)."""
    for bad in ("accel", "=1"):
        out = subprocess.run([exe, os.path.join(scenes_dir, "dingdong.yml"), "--option", bad], capture_output=True, text=True)
        assert out.returncode != 0 and "--option name=value" in out.stderr
