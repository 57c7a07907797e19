"""CPU: the C-ABI shared library loads and exports every symbol include/rt_b200.h declares; without a GPU the
compute entry points fail loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(pkg):
    from cuda_ray_tracer_b200 import binding, build
    build.build()
    return binding.load_library()


def test_header_symbols_exported(pkg, lib):
    from cuda_ray_tracer_b200 import binding
    header = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    declared = sorted(set(re.findall(r"^(?:int|const char \*)\s*(rt_b200_[a-z_0-9]+)\s*\(", header, re.M)))
    assert declared == sorted(binding.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_no_oracle_in_product(pkg):
    """The product must never route through the CPU checkers."""
    pkg_dir = os.path.join(ROOT, "cuda-ray-tracer_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle_py" not in text and "librt_oracle" not in text and "libref_cpu" not in text, f
    import subprocess
    from cuda_ray_tracer_b200 import binding
    needed = subprocess.run(["ldd", binding.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in needed and "ref_cpu" not in needed


def test_version_and_strip_layout(pkg, lib):
    from cuda_ray_tracer_b200 import binding
    assert b"sm_100a" in lib.rt_b200_version()
    assert binding.strip_layout(10, 4, 0, 1).tolist() == list(range(10))
    for h, s, w in [(20, 4, 2), (2160, 8, 8), (450, 8, 4), (7, 3, 2), (5, 8, 4)]:
        seen = []
        sizes = set()
        for r in range(w):
            rows = binding.strip_layout(h, s, r, w)
            sizes.add(len(rows))
            seen.extend(int(y) for y in rows if y >= 0)
        assert len(sizes) == 1                      # equal contribution per rank (all-gather requirement)
        assert sorted(seen) == list(range(h))       # every image row owned exactly once


def test_argument_errors_and_no_gpu(pkg, lib):
    from cuda_ray_tracer_b200 import binding
    from cuda_ray_tracer_b200 import scene as S
    h = C.c_void_p()
    bg = np.zeros(3, dtype=np.float32)
    rc = lib.rt_b200_upload_scene(C.byref(h), 0, 10, 1.0, bg.ctypes.data_as(C.c_void_p), 5, None, 0, None, 0, 0, 0, 1)
    assert rc == 1 and b"empty frame" in lib.rt_b200_last_error()
    rc = lib.rt_b200_upload_scene(C.byref(h), 8, 8, 1.0, bg.ctypes.data_as(C.c_void_p), 5, None, 0, None, 0, 0, 2, 2)
    assert rc == 1 and b"bad rank" in lib.rt_b200_last_error()
    assert lib.rt_b200_update(None, None, None) == 1
    assert lib.rt_b200_gather(None, None, 0, None) == 1
    assert lib.rt_b200_destroy(None) == 0
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        sc = S.Scene(8, 8, 1.0, bg, 5)
        with pytest.raises(binding.RtError, match="rt_b200 error 2"):   # RT_B200_ERR_CUDA: fails loudly, no fallback
            binding.Renderer(sc)
        with pytest.raises(binding.RtError, match="before init_update"):
            binding.update(np.eye(4).reshape(16))


def test_shared_host_block_arguments(pkg, lib, tmp_path):
    """rt_b200_host_alloc_shared: argument checks, and without a GPU the page-lock fails loudly and leaves no file."""
    from cuda_ray_tracer_b200 import binding
    p = C.c_void_p()
    assert lib.rt_b200_host_alloc_shared(None, 4096, 1, C.byref(p)) == 1
    assert lib.rt_b200_host_alloc_shared(b"/nonexistent-dir/x", 4096, 1, C.byref(p)) == 1
    assert b"cannot create" in lib.rt_b200_last_error()
    missing = str(tmp_path / "missing").encode()
    assert lib.rt_b200_host_alloc_shared(missing, 4096, 0, C.byref(p)) == 1 and b"cannot open" in lib.rt_b200_last_error()
    assert lib.rt_b200_host_free_shared(None) == 0
    x = C.c_int(0)
    assert lib.rt_b200_host_free_shared(C.byref(x)) == 1
    path = tmp_path / "frame"
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        ptr = binding.host_alloc_shared(str(path), 10000, True)
        with pytest.raises(binding.RtError, match="cannot create"):     # exclusive: a second creator is refused
            binding.host_alloc_shared(str(path), 10000, True)
        other = binding.host_alloc_shared(str(path), 10000, False)
        binding.host_view(ptr, 50, 50)[:] = 7
        assert int(binding.host_view(other, 50, 50).min()) == 7        # same pages
        binding.host_free_shared(other)
        assert path.exists()
        binding.host_free_shared(ptr)
        assert not path.exists()                                         # the creator unlinks
    else:
        with pytest.raises(binding.RtError, match="rt_b200 error 2"):
            binding.host_alloc_shared(str(path), 10000, True)
        assert not path.exists()


def test_band_layout(pkg, lib):
    """Row bands of the streamed host delivery (rt_b200_update_to_host): cover the rows exactly once, boundaries on whole
    units, every band non-empty; equal / shrinking / growing shapes as option band_taper describes them."""
    from cuda_ray_tracer_b200 import binding
    for rows, unit in ((2160, 16), (4320, 16), (1080, 16), (563, 16), (16, 16), (37, 8), (270, 40)):
        units = -(-rows // unit)
        for n in (1, 2, 3, 5, 8, 12, 16):
            if n > units:
                with pytest.raises(binding.RtError):
                    binding.band_layout(rows, n, unit, 0)
                continue
            for taper in (0, 1, 40, 65, 99, 101, 150, 165, 199):
                s = binding.band_layout(rows, n, unit, taper)
                assert s[0] == 0 and s[-1] == rows and len(s) == n + 1
                if taper == 0:       # equal bands rounded up to whole units may cover the rows in fewer than n bands
                    assert (np.diff(s) >= 0).all() and (s[s < rows] % unit == 0).all()
                    continue
                assert (np.diff(s) > 0).all(), (rows, unit, n, taper, s)
                assert (s[:-1] % unit == 0).all()
    eq = np.diff(binding.band_layout(2160, 8, 16, 0))
    assert eq.max() - eq.min() <= 2160 - 7 * 272 and eq[0] == 272
    sh = np.diff(binding.band_layout(2160, 8, 16, 65))
    gr = np.diff(binding.band_layout(2160, 8, 16, 165))
    assert (np.diff(sh) <= 0).all() and sh[0] > 4 * sh[-1] and sh[-1] <= 64          # a small last band
    assert (np.diff(gr[:-1]) >= 0).all() and gr[0] <= 64 and gr[-2] > 4 * gr[0]       # a small first band
    for bad in ((0, 1, 16, 0), (100, 0, 16, 0), (100, 17, 1, 0), (100, 2, 0, 0), (100, 2, 16, 100), (100, 2, 16, 200), (100, 2, 16, -1)):
        with pytest.raises(binding.RtError):
            binding.band_layout(*bad)
