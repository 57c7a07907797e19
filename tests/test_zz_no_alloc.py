"""rt_b200_render_no_alloc (include/rt_b200.h): one frame without any allocation inside the library — its device scratch
is a static __device__ array; collected last on purpose (the file name).

GPU part: through the C ABI, against the existing entry points (rt_b200_upload_scene + rt_b200_update + rt_b200_gather)
on the same scene and camera, two calls in a row (and more: the scratch is shared by every call and keeps the previous
call's contents).  Tolerance: tests/parity.py (+-1 LSB, <= 0.1 % beyond) — and, both paths being the same exact
arithmetic, byte equality."""
import os

import numpy as np
import pytest

from parity import assert_parity


@pytest.fixture(scope="module")
def rt(pkg):
    from cuda_ray_tracer_b200 import binding
    binding.load_library()
    return binding


def _load(pkg, scenes_dir, name, **over):
    return pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")).with_overrides(**over)


def test_scratch_size_query_and_limit(pkg, rt, scenes_dir):
    """Host-only: the size helper, and the clear error for a scene beyond the static scratch (no GPU needed: the size
    check comes before the first CUDA call)."""
    from cuda_ray_tracer_b200 import camera, synth
    sc = _load(pkg, scenes_dir, "20spheres", width=64, height=48)
    need = rt.no_alloc_scratch_needed(sc)
    assert 20 * 300 < need < 64 * 1024
    assert rt.no_alloc_scratch_needed(synth.synthetic_scene(width=64, height=48)) == 1024000   # C5 (4160 objects) just fits
    big = synth.synthetic_scene(n_spheres=6000, width=64, height=48)
    assert rt.no_alloc_scratch_needed(big) > rt.NO_ALLOC_SCRATCH_BYTES
    with pytest.raises(rt.RtError, match="bytes of device scratch"):
        rt.render_no_alloc(big, camera.TEST_POSES["identity"], 16, 64 * 48 * 4)   # never dereferenced
    with pytest.raises(rt.RtError, match="frame needs"):
        rt.render_no_alloc(sc, camera.TEST_POSES["identity"], 16, 64 * 48 * 4 - 1)
    with pytest.raises(rt.RtError, match="16-byte aligned"):
        rt.render_no_alloc(sc, camera.TEST_POSES["identity"], 8, 64 * 48 * 4)


@pytest.mark.gpu
def test_no_alloc_equals_update_two_calls_in_a_row(pkg, rt, scenes_dir):
    import torch
    from cuda_ray_tracer_b200 import camera, synth
    cases = [(_load(pkg, scenes_dir, "dingdong", width=640, height=360, max_reflections=4), "pose_a"),
             (_load(pkg, scenes_dir, "20spheres", width=801, height=453, max_reflections=4), "identity"),   # ragged tiles
             (_load(pkg, scenes_dir, "reflection_test", width=3840, height=2160, max_reflections=4), "initial"),
             (_load(pkg, scenes_dir, "quadratic", width=320, height=180, max_reflections=4), "pose_b"),
             (synth.synthetic_scene(n_spheres=300, n_cubics=3, n_lights=4, width=192, height=108), "identity")]
    want = []
    for sc, pose in cases:
        with rt.Renderer(sc) as r:                       # the existing entry points
            r.update(camera.TEST_POSES[pose])
            want.append(r.gather())
    out = torch.zeros(3840 * 2160 * 4 + 64, dtype=torch.uint8, device="cuda")   # the CALLER allocates the output
    base = out.data_ptr()
    assert base % 16 == 0

    def frame(sc):
        n = sc.px_width * sc.px_height * 4
        return out[:n].cpu().numpy().reshape(sc.px_height, sc.px_width, 4)

    # two calls in a row on the same scene, then scene after scene through the same scratch, twice round
    for rnd in range(2):
        for (sc, pose), ref in zip(cases, want):
            for call in range(2):
                out.fill_(7)
                ms = rt.render_no_alloc(sc, camera.TEST_POSES[pose], base, out.numel())
                torch.cuda.synchronize()
                got = frame(sc)
                assert ms > 0.0
                assert_parity(got, ref, "no-alloc vs update, round {} call {}".format(rnd, call))
                assert np.array_equal(got, ref), (rnd, call, sc.px_width, sc.px_height)
                n = sc.px_width * sc.px_height * 4
                assert bool((out[n:n + 64] == 7).all()), "wrote past the frame"
    # nothing is allocated by a call: free device memory is the same before and after (kernels are loaded by now)
    sc, pose = cases[0]
    rt.render_no_alloc(sc, camera.TEST_POSES[pose], base, out.numel())
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        rt.render_no_alloc(sc, camera.TEST_POSES[pose], base, out.numel())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    # a host pointer as the output is refused, not dereferenced
    host = np.zeros(sc.px_width * sc.px_height * 4 + 16, dtype=np.uint8)
    addr = (host.ctypes.data + 15) & ~15
    with pytest.raises(rt.RtError, match="not a device pointer"):
        rt.render_no_alloc(sc, camera.TEST_POSES[pose], addr, sc.px_width * sc.px_height * 4)
