"""GPU (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle.

Tolerance (north_star / SURVEY.md §8c): every 8-bit channel within +-1 LSB; at most 0.1 % of pixels may differ
more.  The exact FP64 kernel (option kernel=0) is additionally expected to be almost byte-identical."""
import json
import os

import numpy as np
import pytest

from parity import assert_parity, compare_frames

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
KERNELS = [0, 1]


@pytest.fixture(scope="module")
def rt(pkg):
    from cuda_ray_tracer_b200 import binding
    binding.load_library()          # raises if the CUDA library is missing: no fallback
    return binding


def _load(pkg, scenes_dir, name, **over):
    sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml"))
    return sc.with_overrides(**over) if over else sc


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("name", SCENES)
def test_bundled_scenes_native_resolution(pkg, rt, oracle, scenes_dir, name, kernel):
    from cuda_ray_tracer_b200 import camera
    sc = _load(pkg, scenes_dir, name)
    port = oracle.PortOracle(sc)
    with rt.Renderer(sc) as r:
        r.set_option("kernel", kernel)
        for pose in ("identity", "pose_a", "pose_b"):
            m = camera.TEST_POSES[pose]
            ref_rgb, _, cnt = port.render_rows(m, threads=os.cpu_count(), count=True)
            # default: shadow rays towards lights whose Lambert factor is 0 at the hit are not traced (kernel 1)
            r.set_option("cull_backfacing", 1)
            r.update(m)
            img = r.gather()
            culled = r.counters()
            mask_fn = (lambda sc=sc, m=m: oracle.ill_conditioned_mask(sc, m, os.cpu_count())) if oracle.have_ref() else None
            res = assert_parity(img, oracle.quantize_rgba8(ref_rgb), "{}/{}".format(name, pose), mask_fn=mask_fn)
            assert res["exact_frac"] > 0.999, res
            # ... which cannot change a pixel; with it off the kernel executes the reference's ray census
            r.set_option("cull_backfacing", 0)
            r.update(m)
            assert np.array_equal(r.gather(), img), (name, pose, "culling changed the frame")
            c = r.counters()
            assert culled["shadow_rays"] <= c["shadow_rays"] and culled["lit"] <= c["lit"] and culled["hits"] == c["hits"]
            if res["outliers"] == 0 and res["exact_frac"] == 1.0:
                # same decisions everywhere -> same ray census as the reference would have executed
                for k_gpu, k_cpu in (("primary_rays", "primary_rays"), ("reflection_rays", "reflection_rays"),
                                     ("shadow_rays", "shadow_rays"), ("hits", "hits"), ("lit", "lit")):
                    assert c[k_gpu] == cnt[k_cpu], (k_gpu, c, cnt)
            assert c["primary_rays"] == sc.px_width * sc.px_height
            assert c["shadow_rays"] == c["hits"] * len(sc.lights)


@pytest.mark.parametrize("kernel", KERNELS)
def test_golden_small_frames(pkg, rt, scenes_dir, kernel):
    """Against fixtures minted from the unmodified reference CPU back-end (tests/golden/make_golden.py)."""
    from cuda_ray_tracer_b200 import camera
    meta = json.load(open(os.path.join(HERE, "golden", "golden.json")))["small"]
    frames = np.load(os.path.join(HERE, "golden", "frames_160x120.npz"))
    for name in SCENES:
        sc = _load(pkg, scenes_dir, name, width=meta["width"], height=meta["height"],
                   max_reflections=meta["max_reflections"])
        with rt.Renderer(sc) as r:
            r.set_option("kernel", kernel)
            for pose in meta["poses"]:
                r.update(camera.TEST_POSES[pose])
                assert_parity(r.gather(), frames["{}/{}".format(name, pose)], "{}/{}".format(name, pose))


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("name", ["reflection_test", "dingdong", "cayley"])
def test_full_size_properties_4k(pkg, rt, oracle, scenes_dir, name, kernel):
    """BASELINE.json sizes (3840x2160, 4 bounces): the full CPU frame takes too long for a test, so check a
    sparse oracle (random pixels + whole rows) and size-independent properties."""
    from cuda_ray_tracer_b200 import camera
    sc = _load(pkg, scenes_dir, name, width=3840, height=2160, max_reflections=4)
    m = camera.TEST_POSES["pose_a"]
    port = oracle.PortOracle(sc)
    with rt.Renderer(sc) as r:
        r.set_option("kernel", kernel)
        r.update(m)
        a = r.gather()
        r.set_option("cull_backfacing", 0)                                # trace every shadow ray the reference casts
        r.update(m)
        assert np.array_equal(a, r.gather())                              # deterministic; culling changes no pixel
        c = r.counters()
        assert c["primary_rays"] == 3840 * 2160 and c["shadow_rays"] == c["hits"] * len(sc.lights)
        rng = np.random.default_rng(1)
        xy = np.stack([rng.integers(0, 3840, 20000), rng.integers(0, 2160, 20000)], axis=1).astype(np.int32)
        ref = oracle.quantize_rgba8(port.render_pixels(m, xy))
        assert_parity(a[xy[:, 1], xy[:, 0]][None], ref[None], name + " sparse pixels")
        rows = [0, 1079, 1080, 2159]
        for y in rows:
            rgb, _ = port.render_rows(m, y, y + 1, threads=1)
            assert_parity(a[y:y + 1], oracle.quantize_rgba8(rgb), "{} row {}".format(name, y))
    # strip split: two ranks' rows merged == the single-GPU frame, byte for byte
    merged = np.zeros_like(a)
    for rank in range(2):
        with rt.Renderer(sc, rank=rank, world_size=2) as r:
            r.set_option("kernel", kernel)
            r.set_option("strip_rows", 8)
            r.update(m)
            part = r.gather()
            own = rt.strip_layout(2160, 8, rank, 2)
            own = own[own >= 0]
            merged[own] = part[own]
            other = np.setdiff1d(np.arange(2160), own)
            assert not part[other].any()
    assert np.array_equal(merged, a)


@pytest.mark.parametrize("kernel", KERNELS)
def test_edge_cases(pkg, rt, oracle, kernel):
    from cuda_ray_tracer_b200 import camera
    from cuda_ray_tracer_b200 import scene as S
    bg = np.array([0.25, 0.5, 0.75], dtype=np.float32)

    def check(sc, poses=("identity", "pose_b")):
        port = oracle.PortOracle(sc)
        with rt.Renderer(sc) as r:
            r.set_option("kernel", kernel)
            for pose in poses:
                m = camera.TEST_POSES[pose]
                r.update(m)
                img = r.gather()
                res = compare_frames(img, oracle.quantize_rgba8(port.render(m)))
                assert res["outliers"] <= max(1, int(1e-3 * res["pixels"])) and res["max_diff"] <= 255, res
                if res["pixels"] >= 1000:
                    assert res["outlier_frac"] <= 1e-3, res
        return img

    # empty scene: every pixel is the background
    img = check(S.Scene(33, 17, np.radians(60.0), bg, 5))
    assert (img == oracle.quantize_rgba8(bg.reshape(1, 3))[0]).all()
    # 1x1 frame, and sizes that are not multiples of the tile
    obj = lambda k, refl=0.0, col=(0.8, 0.6, 0.2): S.make_object(k, refl, col)
    lights = np.array([S.light_directional(2.0, [0.3, -1, 0.5], [1, 1, 1]),
                       S.light_spherical(300.0, [2, 5, 3], [1, 0.5, 0.25])], dtype=S.LIGHT_DTYPE)
    for w, h in [(1, 1), (37, 19), (130, 9), (8, 64)]:
        sc = S.Scene(w, h, np.radians(50.0), bg, 3)
        sc.objects = np.array([obj(S.surface_sphere([0, 0, 8], 2.0)), obj(S.surface_plane([0, -2, 0], [0, 1, 0]), 0.4)],
                              dtype=S.OBJECT_DTYPE)
        sc.lights = lights
        check(sc)
    # objects but no lights: hits are black; reflective with max_reflections 0
    sc = S.Scene(64, 48, np.radians(50.0), bg, 0)
    sc.objects = np.array([obj(S.surface_sphere([0, 0, 8], 2.0), 0.5)], dtype=S.OBJECT_DTYPE)
    check(sc)
    # two facing mirrors + a sphere: deep reflection chains, up to max_reflections = 7
    sc = S.Scene(96, 64, np.radians(70.0), bg, 7)
    sc.objects = np.array([obj(S.surface_plane([0, -2, 0], [0, 1, 0]), 0.8), obj(S.surface_plane([0, 3, 0], [0, -1, 0]), 0.7),
                           obj(S.surface_sphere([0.5, 0, 7], 1.0), 0.5, (0.9, 0.1, 0.1))], dtype=S.OBJECT_DTYPE)
    sc.lights = lights
    check(sc)
    # every structural class at once, shuffled order (class sort must keep the lowest-index tie rule),
    # including a duplicated object (exact tie in t) with different colours
    k_quad = np.zeros(20); k_quad[10] = 0.1; k_quad[12] = 0.1; k_quad[17] = 1.0; k_quad[19] = 20.0
    k_cub = S.surface_ding_dong([0, 0, 10])
    sc = S.Scene(120, 90, np.radians(60.0), bg, 4)
    sc.objects = np.array([obj(k_cub), obj(S.surface_sphere([3, 1, 9], 1.0), 0.3), obj(k_quad),
                           obj(S.surface_plane([0, -30, 0], [0, 1, 0]), 0.0, (0.1, 0.5, 0.1)),
                           obj(S.surface_sphere([3, 1, 9], 1.0), 0.0, (0.1, 0.1, 0.9)),
                           obj(S.surface_sphere([-3, 1, 9], 1.5), 0.2)], dtype=S.OBJECT_DTYPE)
    sc.lights = lights
    check(sc, poses=("identity", "pose_a", "pose_b"))
    # point lights sitting (almost) on the surface at the central pixel's hit point (0, 0, 6): the float-rounded shadow
    # direction is shorter than sqrt(1e-7), so the reference's intersect_ray leaves its quadratic branch
    # (surface_impl.h:138) and the FP32 sphere filter, which reasons about the quadratic, must stand aside
    sc = S.Scene(65, 49, np.radians(0.05), bg, 2)
    sc.objects = np.array([obj(S.surface_sphere([0, 0, 8], 2.0), 0.2), obj(S.surface_sphere([0.0003, 0, 6.5], 0.4999), 0.0, (0.2, 0.9, 0.3)),
                           obj(S.surface_sphere([0, 0.5, 6.2], 0.5376), 0.0, (0.2, 0.3, 0.9))] +
                          [obj(S.surface_sphere([4 + i, 3, 9], 0.5)) for i in range(8)], dtype=S.OBJECT_DTYPE)
    sc.lights = np.array([S.light_spherical(1e-4, [1e-5, 0, 6.0001], [1, 1, 1]), S.light_spherical(1e-4, [0, 2e-5, 5.9999], [1, 0.5, 1]),
                          S.light_spherical(1e-4, [-1e-4, 1e-4, 6.0002], [0.5, 1, 1]), S.light_spherical(30.0, [0, 1.0, 5.5], [1, 1, 0.5])],
                         dtype=S.LIGHT_DTYPE)
    check(sc, poses=("identity",))


@pytest.mark.parametrize("kernel", KERNELS)
def test_update_to_host_streams_the_same_frame(pkg, rt, scenes_dir, kernel):
    """rt_b200_update_to_host (banded render, overlapped D2H) delivers exactly the frame update + gather does."""
    import torch
    from cuda_ray_tracer_b200 import camera
    for name, w, h in (("reflection_test", 1920, 1080), ("dingdong", 1001, 563)):
        sc = _load(pkg, scenes_dir, name, width=w, height=h, max_reflections=4)
        host = torch.zeros(w * h * 4, dtype=torch.uint8).pin_memory()
        with rt.Renderer(sc) as r:
            r.set_option("kernel", kernel)
            r.set_option("cull_backfacing", 0)
            r.update(camera.TEST_POSES["pose_a"])
            ref = r.gather()
            c_ref = r.counters()
            for bands, taper in ((0, 0), (1, 0), (3, 0), (16, 0), (8, 70), (5, 40), (16, 99), (0, 65), (7, 150), (16, 199), (0, -1), (0, -1), (0, -1), (0, -1)):
                r.set_option("host_bands", bands)
                r.set_option("band_taper", taper)      # geometrically shrinking bands: same pixels, other boundaries
                r.set_option("band_overlap", bands % 2)  # bands alternating between two streams, or all on one
                host.zero_()
                ms = r.update_to_host(camera.TEST_POSES["pose_a"], host.data_ptr(), host.numel())
                assert ms > 0.0
                assert np.array_equal(host.numpy().reshape(h, w, 4), ref), (name, bands, taper)
                c = r.counters()      # the ray census is band-independent (test counts depend on the shadow cache)
                assert all(c[k] == c_ref[k] for k in ("primary_rays", "reflection_rays", "shadow_rays", "hits", "lit"))
            with pytest.raises(rt.RtError):
                r.update_to_host(camera.TEST_POSES["pose_a"], host.data_ptr(), host.numel() - 1)


def test_update_h_call_surface(pkg, rt, oracle, scenes_dir):
    """init_update / update / cleanup_update with one global context, as the reference's update.h:6-8."""
    from cuda_ray_tracer_b200 import camera
    sc = _load(pkg, scenes_dir, "monkey_saddle", width=200, height=150)
    with pytest.raises(rt.RtError):
        rt.update(camera.IDENTITY)
    rt.init_update(0, sc)
    with pytest.raises(rt.RtError, match="before the first"):
        rt.gather()
    ms = rt.update(camera.TEST_POSES["initial"])
    assert ms > 0.0
    assert_parity(rt.gather(), oracle.quantize_rgba8(oracle.PortOracle(sc).render(camera.TEST_POSES["initial"])))
    rt.cleanup_update()
    rt.cleanup_update()
    with pytest.raises(rt.RtError):
        rt.update(camera.IDENTITY)


def test_reference_build_agrees_on_gpu_box(pkg, rt, oracle, scenes_dir):
    """The prebuilt reference CPU library travels to the GPU box: GPU vs the reference itself (not the port)."""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/libref_cpu.so not present")
    from cuda_ray_tracer_b200 import camera
    for name in ("cayley", "dingdong", "20spheres"):
        sc = _load(pkg, scenes_dir, name, width=400, height=300, max_reflections=4)
        ref = oracle.RefCpu(sc)
        rgb, _ = ref.update(camera.TEST_POSES["pose_a"])
        with rt.Renderer(sc) as r:
            r.update(camera.TEST_POSES["pose_a"])
            assert_parity(r.gather(), oracle.quantize_rgba8(rgb), name)


def _shared_host_worker(rank, world, path, name, w, h, strip, out_dir):
    """One rank of test_shared_host_frame_ranks_as_processes (both ranks use cuda:0: the path needs no peer GPU)."""
    import sys
    root = os.path.dirname(HERE)
    for p in (root, os.path.join(root, "tools")):
        sys.path.insert(0, p)
    from loader import load_package
    pkg = load_package()
    from cuda_ray_tracer_b200 import binding, camera
    sc = pkg.load_from_file(os.path.join(root, "scenes", name + ".yml")).with_overrides(width=w, height=h, max_reflections=4)
    ptr = binding.host_alloc_shared(path, w * h * 4, False)
    ok = True
    with binding.Renderer(sc, rank=rank, world_size=world) as r:
        r.set_option("strip_rows", strip)
        for bands in (0, 1, 2):
            r.set_option("host_bands", bands)
            for pose in ("pose_b", "pose_a"):
                r.update_to_host(camera.TEST_POSES[pose], ptr, w * h * 4)
                if rank == 0:      # complete when rank 0 returns; nobody overwrites it before rank 0's next call
                    got = binding.host_view(ptr, h, w).copy()
                    ok &= bool(np.array_equal(got, np.load(os.path.join(out_dir, pose + ".npy"))))
    binding.host_free_shared(ptr)
    with open(os.path.join(out_dir, "rank{}.txt".format(rank)), "w") as f:
        f.write("ok" if ok else "mismatch")


@pytest.mark.parametrize("world,strip", [(2, 8), (3, 5)])
def test_shared_host_frame_ranks_as_processes(pkg, rt, scenes_dir, tmp_path, world, strip):
    """Multi-rank rt_b200_update_to_host into one shared host frame: every rank (a process) copies its own strips to
    their image rows; rank 0 sees the complete frame, byte-identical to the single-rank frame — ragged last strip,
    padding strips and several bands included."""
    import torch.multiprocessing as mp
    from cuda_ray_tracer_b200 import camera
    name, w, h = "reflection_test", 333, 203
    sc = _load(pkg, scenes_dir, name, width=w, height=h, max_reflections=4)
    with rt.Renderer(sc) as r:
        for pose in ("pose_a", "pose_b"):
            r.update(camera.TEST_POSES[pose])
            np.save(tmp_path / (pose + ".npy"), r.gather())
    path = rt.shared_frame_path(w * h * 4)
    ptr = rt.host_alloc_shared(path, w * h * 4, True)
    try:
        ctx = mp.get_context("spawn")
        procs = [ctx.Process(target=_shared_host_worker, args=(k, world, path, name, w, h, strip, str(tmp_path)))
                 for k in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=180)
        assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
        for k in range(world):
            assert (tmp_path / "rank{}.txt".format(k)).read_text() == "ok"
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
        rt.host_free_shared(ptr)


def test_scenes_larger_than_shared_memory(pkg, rt, oracle):
    """The wavefront kernel keeps planes, quadrics, cubics and lights in shared memory; scenes beyond that (thousands of
    lights or quadrics) must still render — the reference takes any count (update-cpu.cpp:48-76) — through the exact
    kernel, and match the oracle."""
    from cuda_ray_tracer_b200 import camera
    from cuda_ray_tracer_b200 import scene as S
    bg = np.array([0.05, 0.1, 0.2], dtype=np.float32)
    rng = np.random.default_rng(3)
    obj = lambda k, refl=0.0, col=(0.8, 0.6, 0.2): S.make_object(k, refl, col)
    m = camera.TEST_POSES["identity"]
    # 6000 lights (240 KB of light records) over two spheres and a mirror plane
    sc = S.Scene(48, 32, np.radians(55.0), bg, 2)
    sc.objects = np.array([obj(S.surface_sphere([0, 0, 8], 2.0), 0.2), obj(S.surface_sphere([2.5, 1, 7], 0.8), 0.0, (0.2, 0.9, 0.3)),
                           obj(S.surface_plane([0, -2.5, 0], [0, 1, 0]), 0.3, (0.5, 0.5, 0.5))], dtype=S.OBJECT_DTYPE)
    ls = []
    for i in range(6000):
        v = rng.normal(size=3)
        ls.append(S.light_spherical(0.2, list(6.0 * v + np.array([0.0, 6.0, 4.0])), [1, 0.9, 0.8]) if i % 2
                  else S.light_directional(0.001, list(v), [0.8, 0.9, 1.0]))
    sc.lights = np.array(ls, dtype=S.LIGHT_DTYPE)
    with rt.Renderer(sc) as r:
        r.update(m)
        assert_parity(r.gather(), oracle.quantize_rgba8(oracle.PortOracle(sc).render(m, threads=os.cpu_count())), "6000 lights")
    # 3500 quadrics (280 KB of coefficients): small ellipsoids on a lattice
    sc = S.Scene(64, 40, np.radians(55.0), bg, 1)
    objs = []
    for i in range(3500):
        c = np.array([(i % 50) - 25.0, ((i // 50) % 10) - 5.0, 30.0 + 2.0 * (i // 500)]) * 0.9
        a, b, cc = rng.uniform(2.0, 6.0, 3)
        k = np.zeros(20)
        k[10], k[11], k[12] = a, b, cc                                   # a x^2 + b y^2 + c z^2 ...
        k[16], k[17], k[18] = -2 * a * c[0], -2 * b * c[1], -2 * cc * c[2]
        k[19] = a * c[0] ** 2 + b * c[1] ** 2 + cc * c[2] ** 2 - 1.0     # ... = 1 around c
        objs.append(obj(k, 0.0, tuple(rng.uniform(0.2, 0.9, 3))))
    sc.objects = np.array(objs, dtype=S.OBJECT_DTYPE)
    sc.lights = np.array([S.light_directional(2.0, [0.3, -1, 0.5], [1, 1, 1]), S.light_spherical(3000.0, [0, 10, 0], [1, 1, 1])], dtype=S.LIGHT_DTYPE)
    with rt.Renderer(sc) as r:
        assert r.class_counts()["quadric"] == 3500
        r.update(m)
        assert_parity(r.gather(), oracle.quantize_rgba8(oracle.PortOracle(sc).render(m, threads=os.cpu_count())), "3500 quadrics")


def test_contexts_on_several_devices_in_one_process(pkg, rt, scenes_dir):
    """One process, one context per device (rt_b200.h allows a thread per GPU): the >48 KB dynamic shared memory
    attribute of the render kernels is per DEVICE, so the second device must get it too; and the same-process fused
    gather (rt_b200_frame_attach_peer): every rank's kernel stores into rank 0's frame, rank 0's update returns when the
    frame is whole.  With fewer GPUs than ranks the ranks share device 0 — same kernels, same hand-shake."""
    import threading
    import torch
    from cuda_ray_tracer_b200 import camera
    n_dev = torch.cuda.device_count()
    m = camera.TEST_POSES["pose_a"]
    for name in ("dingdong", "20spheres"):
        sc = _load(pkg, scenes_dir, name, width=777, height=431, max_reflections=4)
        with rt.Renderer(sc, device=0) as r:
            r.update(m)
            want = r.gather()
        if n_dev >= 2:      # independent whole frames on two devices from one process
            with rt.Renderer(sc, device=0) as a, rt.Renderer(sc, device=1) as b:
                a.update(m)
                b.update(m)
                assert np.array_equal(a.gather(), want) and np.array_equal(b.gather(), want)
        for world, dynamic in ((2, 1), (4, 1), (3, 0)):
            rs = [rt.Renderer(sc, device=(k if k < n_dev else 0), rank=k, world_size=world) for k in range(world)]
            try:
                for r in rs:
                    r.set_option("strip_rows", 8)
                    r.set_option("dynamic_tiles", dynamic)   # 1: tiles of the whole image from rank 0's counter; 0: own strips
                rs[0].frame_export()
                for r in rs:
                    r.frame_attach_peer(rs[0])
                for frame in range(3):
                    errs = []

                    def run(r):
                        try:
                            r.update(m)
                        except Exception as e:      # noqa: BLE001
                            errs.append(e)
                    ts = [threading.Thread(target=run, args=(r,)) for r in rs[1:]]
                    for t in ts:
                        t.start()
                    rs[0].update(m)                 # returns when every rank's pixels are in rank 0's frame
                    got = rs[0].gather()
                    for t in ts:
                        t.join()
                    assert not errs, errs
                    assert np.array_equal(got, want), (name, world, dynamic, frame)
            finally:
                for r in reversed(rs):
                    r.close()


def test_cubic_filter_and_culling_paths(pkg, rt, oracle):
    """Kernel 1's FP32 cubic filter (cached halves / generic path / many lights) and the light culling: with every
    combination of the two options the frame equals the exact kernel's, and the oracle's."""
    from cuda_ray_tracer_b200 import camera
    from cuda_ray_tracer_b200 import scene as S
    bg = np.array([0.1, 0.2, 0.3], dtype=np.float32)
    rng = np.random.default_rng(5)

    def lights(n):
        out = []
        for i in range(n):
            v = rng.normal(size=3)
            if i % 3 == 0:
                out.append(S.light_spherical(200.0 / n + 20.0, list(4.0 * v + np.array([0.0, 3.0, 4.0])), [1, 0.9, 0.8]))
            else:
                out.append(S.light_directional(3.0 / n + 0.2, list(v), [0.8, 0.9, 1.0]))
        return np.array(out, dtype=S.LIGHT_DTYPE)

    ding = S.surface_polynomial({"y3": 1.0, "x2": 1.0, "y2": -1.0, "z2": 1.0, "z": -20.0, "c": 100.0})
    cases = {
        "one cubic, cached halves": ([S.make_object(S.surface_cayley(), 0.3, (0.8, 0.7, 0.6))], lights(5)),
        "one cubic, 300 lights (no per-light cache, 10 light chunks)": ([S.make_object(ding, 0.0, (0.7, 0.7, 0.9))], lights(300)),
        "two cubics and a sphere": ([S.make_object(S.surface_clebsch(), 0.0, (0.9, 0.5, 0.5)),
                                     S.make_object(ding, 0.4, (0.5, 0.9, 0.5)),
                                     S.make_object(S.surface_sphere([0.5, 0.5, 6.0], 1.0), 0.2, (0.5, 0.5, 0.9))], lights(7)),
    }
    for label, (objs, ls) in cases.items():
        sc = S.Scene(96, 54, np.radians(50.0), bg, 3)
        sc.objects = np.array(objs, dtype=S.OBJECT_DTYPE)
        sc.lights = ls
        port = oracle.PortOracle(sc)
        with rt.Renderer(sc) as r:
            for pose in ("identity", "pose_a"):
                m = camera.TEST_POSES[pose]
                r.set_option("kernel", 0)
                r.update(m)
                exact = r.gather()
                assert_parity(exact, oracle.quantize_rgba8(port.render(m, threads=os.cpu_count())), label + " exact kernel")
                r.set_option("kernel", 1)
                seen = {}
                for filt in (0, 1):
                    for cull in (0, 1):
                        r.set_option("cubic_filter", filt)
                        r.set_option("cull_backfacing", cull)
                        r.update(m)
                        assert np.array_equal(r.gather(), exact), (label, pose, filt, cull)
                        seen[(filt, cull)] = r.counters()
                assert seen[(1, 0)]["cubic_exact"] <= seen[(0, 0)]["cubic_exact"]
                assert seen[(1, 1)]["shadow_rays"] <= seen[(1, 0)]["shadow_rays"]
                assert seen[(0, 0)]["shadow_rays"] == seen[(0, 0)]["hits"] * len(ls)


def test_sphere_grid_frames_identical(pkg, rt, oracle, scenes_dir):
    """Option accel (csrc/rt_grid.cuh): with the sphere class walked through the uniform grid the frame
    is byte-identical to the brute-force loops (wavefront kernel with accel = 0, and the exact kernel), at every grid
    density and in both kernel variants; the ray census is unchanged and far fewer spheres are examined."""
    from cuda_ray_tracer_b200 import camera, synth
    from cuda_ray_tracer_b200 import scene as S
    bg = np.array([0.2, 0.3, 0.4], dtype=np.float32)
    rng = np.random.default_rng(11)
    obj = lambda k, refl=0.0, col=(0.8, 0.6, 0.2): S.make_object(k, refl, col)

    def cloud(n, scale, extras):
        objs = []
        for i in range(n):
            c = rng.uniform(-1, 1, 3) * scale * np.array([1.0, 0.6, 1.0]) + np.array([0, 0, 2.2 * scale])
            objs.append(obj(S.surface_sphere(list(c), float(rng.uniform(0.02, 0.09) * scale)), 0.3 if i % 4 == 0 else 0.0,
                            tuple(rng.uniform(0.1, 0.9, 3))))
        return objs + extras

    far_pose = camera.camera_matrix((3.0e5, 2.0, -5.0), 80.0, -10.0)      # the walker refuses such origins: brute force
    lights = np.array([S.light_spherical(3000.0, [3, 9, 14], [1, 0.9, 0.8]), S.light_spherical(2000.0, [-6, -2, 25], [0.8, 0.9, 1]),
                       S.light_directional(1.5, [0.3, -1, 0.5], [1, 1, 1])], dtype=S.LIGHT_DTYPE)
    dup = S.surface_sphere([0.5, 0.25, 20.0], 1.5)
    cases = {}
    sc = synth.synthetic_scene(n_spheres=600, n_cubics=3, n_lights=4, width=192, height=108)
    cases["synthetic 600 spheres + 3 cubics"] = (sc, ("identity", "pose_a"))
    sc = S.Scene(160, 100, np.radians(60.0), bg, 4)
    sc.objects = np.array(cloud(300, 10.0, [
        obj(dup, 0.0, (0.9, 0.1, 0.1)), obj(dup, 0.5, (0.1, 0.9, 0.1)),                  # exact tie: lowest index wins
        obj(S.surface_sphere([0.5, 0.25, 20.0], 0.7), 0.0),                              # nested
        obj(S.surface_sphere([0, 0, 22.0], 400.0), 0.0, (0.3, 0.3, 0.5)),                # huge, contains the camera: loose
        obj(S.surface_sphere([1, 1, 15.0], 1e-6), 0.0),                                  # tiny: loose
        obj(S.surface_sphere([500.0, 0, 22.0], 3.0), 0.0),                               # outlier stretching the bounds
        obj(S.surface_plane([0, -7, 0], [0, 1, 0]), 0.4, (0.5, 0.5, 0.5))]), dtype=S.OBJECT_DTYPE)
    sc.lights = lights
    cases["random cloud with awkward spheres + mirror plane"] = (sc, ("identity", "pose_b", far_pose))
    cases["20spheres (grid forced on)"] = (_load(pkg, scenes_dir, "20spheres"), ("identity", "pose_a"))
    # point lights on the surface (see test_edge_cases): shadow directions below the quadratic-branch threshold
    sc = S.Scene(65, 49, np.radians(0.05), bg, 2)
    sc.objects = np.array([obj(S.surface_sphere([0, 0, 8], 2.0), 0.2), obj(S.surface_sphere([0.0003, 0, 6.5], 0.4999), 0.0, (0.2, 0.9, 0.3)),
                           obj(S.surface_sphere([0, 0.5, 6.2], 0.5376), 0.0, (0.2, 0.3, 0.9))] +
                          [obj(S.surface_sphere([4 + i, 3, 9], 0.5)) for i in range(8)], dtype=S.OBJECT_DTYPE)
    sc.lights = np.array([S.light_spherical(1e-4, [1e-5, 0, 6.0001], [1, 1, 1]), S.light_spherical(1e-4, [0, 2e-5, 5.9999], [1, 0.5, 1]),
                          S.light_spherical(30.0, [0, 1.0, 5.5], [1, 1, 0.5])], dtype=S.LIGHT_DTYPE)
    cases["lights on the surface"] = (sc, ("identity",))

    for label, (sc, poses) in cases.items():
        port = oracle.PortOracle(sc)
        with rt.Renderer(sc) as r:
            info = r.grid_info()
            assert info["gridded"] + info["loose"] == r.class_counts()["sphere"], info
            for pose in poses:
                m = camera.TEST_POSES[pose] if isinstance(pose, str) else pose
                r.set_option("kernel", 0)
                r.update(m)
                exact = r.gather()
                if isinstance(pose, str):
                    assert_parity(exact, oracle.quantize_rgba8(port.render(m, threads=os.cpu_count())), label + " exact kernel")
                r.set_option("kernel", 1)
                r.set_option("cull_backfacing", 0)
                r.set_option("accel", 0)
                r.set_option("grid_density", 0)
                r.update(m)
                assert np.array_equal(r.gather(), exact), (label, pose, "brute force")
                brute = r.counters()
                for variant, density in ((0, 0), (3, 0), (0, 20), (0, 900)):
                    r.set_option("accel", 1)
                    r.set_option("variant", variant)
                    r.set_option("grid_density", density)
                    r.update(m)
                    assert np.array_equal(r.gather(), exact), (label, pose, variant, density, r.grid_info())
                    c = r.counters()
                    for k in ("primary_rays", "reflection_rays", "shadow_rays", "hits", "lit"):
                        assert c[k] == brute[k], (label, pose, k, c, brute)
                    if label.startswith("synthetic") and density == 0:
                        assert c["tests_sphere"] < 0.25 * brute["tests_sphere"], (c, brute)
                r.set_option("variant", 0)
                r.set_option("cull_backfacing", 1)
            # the streamed / banded entry point with the grid on
            import torch
            host = torch.zeros(sc.px_width * sc.px_height * 4, dtype=torch.uint8).pin_memory()
            r.update_to_host(m, host.data_ptr(), host.numel())
            assert np.array_equal(host.numpy().reshape(sc.px_height, sc.px_width, 4), exact), (label, "update_to_host")
