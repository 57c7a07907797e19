"""GPU fuzzing of kernel 1 (FP32 cubic filter, light culling, caches) against kernel 0 (exact, thread per pixel):
random cubic-class surfaces, random lights and cameras, small frames; every frame must be byte-identical.
    python tools/gpu_fuzz_scenes.py [n_scenes] [seed] [max_cubics] [mixed]
mixed = 1: also planes (some mirrors), quadrics and up to 12 spheres per scene, and scenes without any cubic — the plain kernel
instantiation, the per-light shadow cache over all classes, the sphere-first / cubic-first search order, exact ties."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera
from cuda_ray_tracer_b200 import scene as S

n_scenes = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
max_cubics = int(sys.argv[3]) if len(sys.argv) > 3 else 3
mixed = len(sys.argv) > 4 and sys.argv[4] == "1"
bad = 0
undecided = tests = 0
_last_sphere = None
for it in range(n_scenes):
    objs = []
    for _ in range(rng.integers(0 if mixed else 1, max_cubics + 1)):
        k = np.zeros(20)
        style = rng.integers(0, 4)
        if style == 0:
            k = rng.normal(size=20) * 10.0 ** rng.uniform(-1, 1)
        elif style == 1:
            idx = rng.choice(20, size=rng.integers(2, 7), replace=False)
            k[idx] = rng.normal(size=len(idx)) * 10.0 ** rng.uniform(-2, 2, size=len(idx))
            if not k[:10].any():
                k[rng.integers(0, 10)] = rng.normal()
        elif style == 2:     # quadric plus a faint cubic term: the reference's degree / discriminant switches are in play
            k[10:13] = rng.uniform(0.5, 2.0, size=3)
            k[19] = -rng.uniform(1.0, 9.0)
            k[rng.integers(0, 10)] = 10.0 ** rng.uniform(-8, -2)
        else:
            k = [S.surface_cayley(), S.surface_clebsch(), S.surface_ding_dong([0.0, 0.0, 3.0])][rng.integers(0, 3)]
        objs.append(S.make_object(np.asarray(k, dtype=np.float64), float(rng.choice([0.0, 0.0, 0.4])), tuple(rng.uniform(0.2, 1.0, size=3))))
    if rng.random() < 0.3:
        objs.append(S.make_object(S.surface_sphere(list(rng.uniform(-3, 3, size=3) + [0, 0, 6]), rng.uniform(0.3, 1.5)), 0.2, (0.6, 0.6, 0.9)))
    if mixed:
        for _ in range(rng.integers(0, 3)):       # planes, every other one a mirror
            nrm = rng.normal(size=3)
            objs.append(S.make_object(S.surface_plane(list(rng.uniform(-6, 6, size=3)), list(nrm)), float(rng.choice([0.0, 0.5])), tuple(rng.uniform(0.2, 1.0, size=3))))
        for _ in range(rng.integers(0, 13)):      # spheres, sometimes an exact duplicate of the previous one (tie in t)
            if rng.random() < 0.1 and _last_sphere is not None:
                objs.append(S.make_object(_last_sphere, 0.0, tuple(rng.uniform(0.2, 1.0, size=3))))
            else:
                _last_sphere = S.surface_sphere(list(rng.uniform(-5, 5, size=3) + [0, 0, 7]), float(rng.uniform(0.2, 2.0)))
                objs.append(S.make_object(_last_sphere, float(rng.choice([0.0, 0.0, 0.3])), tuple(rng.uniform(0.2, 1.0, size=3))))
        for _ in range(rng.integers(0, 3)):       # quadrics: ellipsoids, paraboloids, hyperboloids around random centres
            k = np.zeros(20)
            k[10:13] = rng.uniform(0.2, 2.0, size=3) * rng.choice([1.0, 1.0, -1.0], size=3)
            k[13:16] = rng.normal(size=3) * rng.choice([0.0, 0.3])
            k[16:19] = rng.normal(size=3) * 2.0
            k[19] = rng.normal() * 4.0
            objs.append(S.make_object(k, float(rng.choice([0.0, 0.2])), tuple(rng.uniform(0.2, 1.0, size=3))))
        if not objs:
            objs.append(S.make_object(S.surface_sphere([0, 0, 6], 1.0), 0.0, (0.5, 0.5, 0.5)))
        order = rng.permutation(len(objs))        # class order must not matter (lowest ORIGINAL index wins ties)
        objs = [objs[i] for i in order]
    ls = []
    for _ in range(rng.integers(1, 8)):
        if rng.random() < 0.4:
            ls.append(S.light_spherical(rng.uniform(20, 400), list(rng.uniform(-8, 8, size=3)), [1, 1, 1]))
        else:
            v = rng.normal(size=3)
            if rng.random() < 0.3:
                v = np.eye(3)[rng.integers(0, 3)] * rng.choice([-1.0, 1.0])
            ls.append(S.light_directional(rng.uniform(0.3, 2.0), list(v), [1, 0.9, 0.8]))
    sc = S.Scene(128, 72, np.radians(rng.uniform(30, 90)), np.array(rng.uniform(0, 1, size=3), dtype=np.float32), int(rng.integers(0, 5)))
    sc.objects = np.array(objs, dtype=S.OBJECT_DTYPE)
    sc.lights = np.array(ls, dtype=S.LIGHT_DTYPE)
    m = camera.camera_matrix(tuple(rng.uniform(-6, 6, size=3)), float(rng.uniform(0, 360)), float(rng.uniform(-60, 60)))
    with binding.Renderer(sc) as r:
        r.set_option("kernel", 0)
        r.update(m)
        exact = r.gather()
        r.set_option("kernel", 1)
        r.update(m)
        img = r.gather()
        c = r.counters()
        tests += c["tests_cubic"]; undecided += c["cubic_exact"]
        if not np.array_equal(img, exact):
            bad += 1
            d = np.abs(img.astype(int) - exact.astype(int)).max(axis=2)
            print("MISMATCH scene", it, "pixels", int((d > 0).sum()), "max", int(d.max()), flush=True)
            np.save("gpurun_out/fuzz_bad_scene_%d.npy" % it, {"objects": sc.objects, "lights": sc.lights, "m": m, "fov": sc.vertical_fov,
                                                             "bg": sc.bg_color, "refl": sc.max_reflections}, allow_pickle=True)
print("GPU FUZZ", "OK" if bad == 0 else "FAILED", n_scenes, "scenes,", bad, "mismatching; cubic tests", tests, "exact", undecided)
sys.exit(1 if bad else 0)
