"""GPU (-m gpu): parity at every configuration BASELINE.json names, at the configuration's own size.

  C2  quadratic + cubic @ 1920x1080          full CPU frame (seconds on the host cores)
  C3  cayley, clebsch, dingdong, monkey_saddle @ 3840x2160, 4 bounces          } sparse oracle: 20 000 random pixels
  C4  reflection_test @ 3840x2160, 4 bounces (+ 20spheres, quadratic, cubic:   }   + 4 whole rows per frame, and the
      bench.py's `bundled4k` step times all eight at this size)                }   frame byte-identical to kernel 0
  C5  4096 spheres + 64 cubics + 16 point lights @ 7680x4320, brute force and sphere grid: 2 000 random pixels + 2
      whole rows through the restatement (a pixel costs ~0.75 ms of CPU), 300 of them also through the unmodified
      reference build, and the frame byte-identical to kernel 0.

The reference hook for a sparse oracle is render_pixel(), src/update-cpu.cpp:82 (external linkage); the acceptance
rule is tests/parity.py (+-1 LSB per channel, <= 0.1 % of pixels beyond)."""
import os

import numpy as np
import pytest

from parity import assert_parity

pytestmark = pytest.mark.gpu

BUNDLED = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]


@pytest.fixture(scope="module")
def rt(pkg):
    from cuda_ray_tracer_b200 import binding
    binding.load_library()          # raises if the CUDA library is missing: no fallback
    return binding


def _load(pkg, scenes_dir, name, **over):
    return pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(**over)


def _sparse_check(oracle, port, ref, m, img, n_random, rows, n_ref, label, seed=7):
    """Random pixels + whole rows of `img` against the restatement (threaded), a subset against the reference build."""
    h, w = img.shape[:2]
    rng = np.random.default_rng(seed)
    xy = np.stack([rng.integers(0, w, n_random), rng.integers(0, h, n_random)], axis=1).astype(np.int32)
    for y in rows:
        xy = np.concatenate([xy, np.stack([np.arange(w), np.full(w, y)], axis=1).astype(np.int32)])
    want = oracle.quantize_rgba8(port.render_pixels_mt(m, xy))
    res = assert_parity(img[xy[:, 1], xy[:, 0]][None], want[None], label + " (restatement)")
    if ref is not None and n_ref:
        sub = xy[:n_ref]
        want_ref = oracle.quantize_rgba8(ref.render_pixels(m, sub))
        assert_parity(img[sub[:, 1], sub[:, 0]][None], want_ref[None], label + " (reference build)")
        assert np.array_equal(want_ref, want[:n_ref]), label + ": restatement and reference build disagree"
    return res


@pytest.mark.parametrize("name", ["quadratic", "cubic"])
def test_c2_full_frames_1080p(pkg, rt, oracle, scenes_dir, name):
    """BASELINE.json configs[1]: the closed-form quadratic / cubic root paths at 1920x1080, whole frame vs the CPU."""
    from cuda_ray_tracer_b200 import camera
    sc = _load(pkg, scenes_dir, name, width=1920, height=1080)
    port = oracle.PortOracle(sc)
    with rt.Renderer(sc) as r:
        for pose in ("initial", "pose_a"):
            m = camera.TEST_POSES[pose]
            want = oracle.quantize_rgba8(port.render(m, threads=os.cpu_count()))
            r.set_option("kernel", 1)
            r.update(m)
            img = r.gather()
            mask_fn = (lambda m=m: oracle.ill_conditioned_mask(sc, m, os.cpu_count())) if oracle.have_ref() else None
            res = assert_parity(img, want, "{} 1080p {}".format(name, pose), mask_fn=mask_fn)
            assert res["exact_frac"] > 0.999, res
            r.set_option("kernel", 0)
            r.update(m)
            assert np.array_equal(r.gather(), img), (name, pose, "kernel 0 differs from kernel 1")
    if oracle.have_ref():       # the unmodified reference build on a band of the same frame
        ref = oracle.RefCpu(sc)
        rgb, _ = ref.render_rows(camera.TEST_POSES["pose_a"], 500, 560, 1, threads=1)
        assert_parity(img[500:560], oracle.quantize_rgba8(rgb), name + " 1080p rows 500..559 (reference build)")


@pytest.mark.parametrize("name", BUNDLED)
def test_c3_c4_all_bundled_scenes_4k(pkg, rt, oracle, scenes_dir, name):
    """BASELINE.json configs[2..3] and every frame bench.py's bundled4k step times: 3840x2160, max_reflections 4."""
    from cuda_ray_tracer_b200 import camera
    sc = _load(pkg, scenes_dir, name, width=3840, height=2160, max_reflections=4)
    port = oracle.PortOracle(sc)
    ref = oracle.RefCpu(sc) if oracle.have_ref() else None
    with rt.Renderer(sc) as r:
        for pose in ("initial", "pose_b"):        # "initial" is the pose bench.py times
            m = camera.TEST_POSES[pose]
            r.set_option("kernel", 1)
            r.update(m)
            img = r.gather()
            c = r.counters()
            assert c["primary_rays"] == 3840 * 2160
            res = _sparse_check(oracle, port, ref, m, img, 20000, [0, 1079, 1080, 2159], 1000,
                                "{} 4K {}".format(name, pose))
            assert res["exact_frac"] > 0.999, res
            r.set_option("kernel", 0)
            r.update(m)
            assert np.array_equal(r.gather(), img), (name, pose, "kernel 0 differs from kernel 1")


@pytest.mark.timeout(900)
def test_c5_synthetic_8k(pkg, rt, oracle):
    """BASELINE.json configs[4] at its own size, brute force (accel 0) and through the sphere grid (accel 1)."""
    from cuda_ray_tracer_b200 import camera, synth
    sc = synth.synthetic_scene()
    assert (sc.px_width, sc.px_height, len(sc.objects), len(sc.lights)) == (7680, 4320, 4160, 16)
    m = camera.TEST_POSES["initial"]
    port = oracle.PortOracle(sc)
    ref = oracle.RefCpu(sc) if oracle.have_ref() else None
    with rt.Renderer(sc) as r:
        frames = {}
        for accel in (0, 1):
            r.set_option("kernel", 1)
            r.set_option("accel", accel)
            r.set_option("cull_backfacing", 0)
            r.update(m)
            frames[accel] = r.gather()
            c = r.counters()
            assert c["primary_rays"] == 7680 * 4320 and c["shadow_rays"] == c["hits"] * 16
        assert np.array_equal(frames[0], frames[1]), "sphere grid changed the frame"
        res = _sparse_check(oracle, port, ref, m, frames[0], 2000, [1500, 2600], 300, "C5 8K")
        assert res["exact_frac"] > 0.999, res
        r.set_option("kernel", 0)       # one thread per pixel over 4160 objects x 16 lights: seconds, once
        r.update(m)
        assert np.array_equal(r.gather(), frames[0]), "kernel 0 differs from kernel 1 on C5"
