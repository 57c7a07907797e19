"""CPU: the oracle restatement (oracle/rt_oracle.c) against the reference's own output.

The reference ships no tests or golden vectors.  The pins are (1) fixtures minted from the
reference's CPU back-end compiled unmodified (tests/golden/make_golden.py), (2) that build itself when it
is present (oracle/_ref), (3) hand-solvable known answers, (4) the survey's non-background fractions."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
SCENES = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
# SURVEY.md / BASELINE.md §2: non-background pixels at native resolution, identity camera (percent, 1 decimal)
SURVEY_NON_BG = {"20spheres": 15.8, "quadratic": 50.0, "cubic": 100.0, "cayley": 42.7, "clebsch": 38.8,
                 "dingdong": 49.1, "monkey_saddle": 50.7, "reflection_test": 50.7}


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(GOLD, "golden.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(GOLD, "frames_160x120.npz")), np.load(os.path.join(GOLD, "rays.npz"))


def _same_bits(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.array_equal(a.view(np.uint64), b.view(np.uint64)) or \
        bool(np.all((a.view(np.uint64) == b.view(np.uint64)) | (np.isnan(a) & np.isnan(b))))


@pytest.mark.parametrize("name", SCENES)
def test_port_matches_golden_small_frames(pkg, oracle, golden, scenes_dir, name):
    meta, frames, _ = golden
    from cuda_ray_tracer_b200 import camera
    sm = meta["small"]
    sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(
        width=sm["width"], height=sm["height"], max_reflections=sm["max_reflections"])
    port = oracle.PortOracle(sc)
    for pose in sm["poses"]:
        got = oracle.quantize_rgba8(port.render(camera.TEST_POSES[pose], threads=4))
        assert np.array_equal(got, frames["{}/{}".format(name, pose)]), (name, pose)


@pytest.mark.parametrize("name", SCENES)
def test_port_native_hash_and_survey_fraction(pkg, oracle, golden, scenes_dir, name):
    meta = golden[0]["scenes"][name]
    sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml"))
    assert (sc.px_width, sc.px_height) == (meta["width"], meta["height"])
    q = oracle.quantize_rgba8(oracle.PortOracle(sc).render(threads=8))
    assert oracle.fnv1a64(q[..., :3].tobytes()) == meta["fnv1a64_rgb_identity"]
    bgq = oracle.quantize_rgba8(sc.bg_color.reshape(1, 3))[0]
    frac = float((q != bgq).any(axis=-1).mean())
    assert abs(frac - meta["non_background_frac"]) < 1e-12
    assert abs(100 * frac - SURVEY_NON_BG[name]) < 0.1


def test_port_matches_golden_rays(oracle, golden):
    rays = golden[2]
    for i, k in enumerate(rays["coefs"]):
        t = np.array([oracle.PortOracle.intersect_ray(k, o, d) for o, d in zip(rays["origins"], rays["dirs"])])
        n = np.array([oracle.PortOracle.normal_vector(k, o) for o in rays["origins"]])
        assert _same_bits(t, rays["t"][i]), i
        assert _same_bits(n, rays["normals"][i]), i


def test_port_matches_reference_build_bitwise(pkg, oracle, scenes_dir):
    """Float-bit equality of whole frames against the unmodified reference, when it is built here."""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    from cuda_ray_tracer_b200 import camera
    for name in SCENES:
        sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(width=200, height=150)
        ref = oracle.RefCpu(sc)
        port = oracle.PortOracle(sc)
        for pose in ("initial", "pose_b"):
            a, _ = ref.update(camera.TEST_POSES[pose])
            b = port.render(camera.TEST_POSES[pose])
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (name, pose)
        xy = np.array([[0, 0], [199, 149], [100, 75], [17, 101]], dtype=np.int32)
        assert np.array_equal(ref.render_pixels(camera.TEST_POSES["pose_a"], xy).view(np.uint32),
                              port.render_pixels(camera.TEST_POSES["pose_a"], xy).view(np.uint32))
        rows_ref, _ = ref.render_rows(camera.IDENTITY, 3, 150, 40, threads=2)
        rows_port, _ = port.render_rows(camera.IDENTITY, 3, 150, 40, threads=3)
        assert np.array_equal(rows_ref.view(np.uint32), rows_port.view(np.uint32))


def test_known_answers(pkg, oracle):
    """Hand-solvable cases of intersect_ray / normal_vector (surface_impl.h:105-172)."""
    from cuda_ray_tracer_b200 import scene as S
    I = oracle.PortOracle.intersect_ray
    sphere = S.surface_sphere([0, 0, 10], 2.0)
    assert I(sphere, [0, 0, 0], [0, 0, 1]) == 8.0                  # near root
    assert I(sphere, [0, 0, 10], [0, 0, 1]) == 2.0                 # inside: near root < EPS -> far root
    assert I(sphere, [0, 5, 0], [0, 0, 1]) == -1.0                 # miss: negative discriminant
    assert I(sphere, [0, 0, 20], [0, 0, 1]) == -8.0                # behind: far root returned, negative
    plane = S.surface_plane([0, -2, 0], [0, 1, 0])
    assert I(plane, [0, 0, 0], [0, -1, 0]) == 2.0
    assert I(plane, [0, 0, 0], [1, 0, 0]) == -1.0                  # parallel: |t1| <= EPS
    # x^3 - 1 = 0 along +x from the origin: single real root (Cardano), t = 1
    k = np.zeros(20); k[0] = 1.0; k[19] = -1.0
    assert abs(I(k, [0, 0, 0], [1, 0, 0]) - 1.0) < 1e-15
    # (x-1)(x-2)(x-3): three real roots (trigonometric branch), smallest >= EPS is 1
    k = np.zeros(20); k[0] = 1.0; k[10] = -6.0; k[16] = 11.0; k[19] = -6.0
    assert abs(I(k, [0, 0, 0], [1, 0, 0]) - 1.0) < 1e-12
    assert abs(I(k, [1.5, 0, 0], [1, 0, 0]) - 0.5) < 1e-12         # from between the roots: next is x = 2
    # quadratic with t2 < 0 returns the LARGER root first (surface_impl.h:146-148 quirk)
    k = np.zeros(20); k[10] = -1.0; k[19] = 4.0                     # 4 - x^2
    assert I(k, [0, 0, 0], [1, 0, 0]) == 2.0
    assert I(k, [-5, 0, 0], [1, 0, 0]) == 7.0                       # (-t1-sqrt)/(2 t2) with t2<0 is the far root
    n = oracle.PortOracle.normal_vector(sphere, [0, 0, 8])
    assert np.allclose(n, [0, 0, -1])                               # gradient direction, not flipped to the viewer
    assert np.allclose(oracle.PortOracle.normal_vector(plane, [3, -2, 1]), [0, 1, 0])


def test_reflection_and_empty_scene(pkg, oracle):
    from cuda_ray_tracer_b200 import scene as S
    bg = np.array([0.25, 0.5, 0.75], dtype=np.float32)
    empty = S.Scene(16, 8, np.radians(60.0), bg, 5)
    img = oracle.PortOracle(empty).render()
    assert np.array_equal(img, np.broadcast_to(bg, (8, 16, 3)))
    # mirror plane, no lights, max_reflections 0: colour = (1-r)*0 + r*bg on the plane
    sc = S.Scene(8, 8, np.radians(60.0), bg, 0)
    sc.objects = np.array([S.make_object(S.surface_plane([0, -1, 0], [0, 1, 0]), 0.5, [1, 1, 1])], dtype=S.OBJECT_DTYPE)
    img = oracle.PortOracle(sc).render()
    assert np.array_equal(img[0, 0], np.float32(0.5) * np.float32(0.0) + np.float32(0.5) * bg)
    assert np.array_equal(img[7, 0], bg)


def test_quantize_rule(oracle):
    c = np.array([[0.0, 0.5, 1.0], [0.1, 0.998, 0.00196]], dtype=np.float32)
    q = oracle.quantize_rgba8(c)
    assert q.tolist() == [[0, 128, 255, 255], [26, 254, 0, 255]]


def test_ill_conditioned_mask(pkg, oracle, scenes_dir):
    """the mask of pixels where the reference's own result flips under a one-rounding perturbation
    (FMA contraction, camera matrix + 1 ulp), and the acceptance rule 'outliers only inside the mask' in tests/parity.py."""
    import numpy as np
    from cuda_ray_tracer_b200 import camera
    from parity import assert_parity
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/libref_cpu.so not present")
    load = lambda n: pkg.load_from_file(os.path.join(scenes_dir, n + ".yml")).with_overrides(width=320, height=240)
    threads = os.cpu_count() or 1
    # spheres, planes and quadrics are well-conditioned everywhere
    assert not oracle.ill_conditioned_mask(load("quadratic"), camera.TEST_POSES["identity"], threads).any()
    assert not oracle.ill_conditioned_mask(load("20spheres"), camera.TEST_POSES["pose_a"], threads).any()
    # cayley seen from ON the surface (every primary ray has t0 = t1 = 0, delta = 0 analytically): speckle
    sc = load("cayley")
    m = camera.TEST_POSES["identity"]
    mask = oracle.ill_conditioned_mask(sc, m, threads)
    assert 0 < mask.sum() < 0.02 * mask.size
    want = oracle.quantize_rgba8(oracle.RefCpu(sc).render_rows(m, threads=threads)[0])
    ys, xs = np.nonzero(mask)
    inside = want.copy()
    inside[ys[:20], xs[:20], :3] ^= 0x40                 # 20 pixels far off, all inside the mask: accepted
    r = assert_parity(inside, want, "inside", mask_fn=lambda: mask)
    assert r["outliers"] == 20 and r["outliers_outside_mask"] == 0
    yo, xo = np.nonzero(~mask)
    outside = want.copy()
    outside[yo[0], xo[0], :3] ^= 0x40                    # one pixel outside it: refused
    with pytest.raises(AssertionError, match="outside the ill-conditioned mask"):
        assert_parity(outside, want, "outside", mask_fn=lambda: mask)
    assert_parity(want, want, "exact", mask_fn=lambda: 1 / 0)   # no outliers: the mask is never computed
