"""CPU: host-side scene model (cuda-ray-tracer_b200/scene.py) against the reference's factories and loader
semantics (src/surface.cpp, src/light.cpp, src/scene.cpp)."""
import os

import numpy as np
import pytest

SCENES = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]


def test_record_layouts(pkg):
    from cuda_ray_tracer_b200 import scene as S
    assert S.OBJECT_DTYPE.itemsize == 176 and S.OBJECT_DTYPE.fields["reflection_ratio"][1] == 160
    assert S.OBJECT_DTYPE.fields["color"][1] == 164
    assert S.LIGHT_DTYPE.itemsize == 48 and S.LIGHT_DTYPE.fields["p"][1] == 8
    assert S.LIGHT_DTYPE.fields["light_color"][1] == 32


def test_factories_match_reference(pkg, oracle):
    """Bitwise, against the reference's own factory code when that build is present."""
    if not oracle.have_ref():
        pytest.skip("oracle/_ref not built")
    from cuda_ray_tracer_b200 import scene as S
    R = oracle.RefCpu
    rng = np.random.default_rng(7)
    for _ in range(20):
        c, n, r = rng.normal(size=3) * 10, rng.normal(size=3), abs(rng.normal()) + 0.1
        assert np.array_equal(S.surface_sphere(c, r), R.make_surface("sphere", c, s=r))
        assert np.array_equal(S.surface_plane(c, n), R.make_surface("plane", c, n))
        assert np.array_equal(S.surface_ding_dong(c), R.make_surface("dingDong", c))
        col = rng.uniform(size=3).astype(np.float32)
        inten = float(np.float32(rng.uniform(0, 900)))
        assert S.light_directional(inten, n, col).tobytes() == R.make_light(False, inten, n, col).tobytes()
        assert S.light_spherical(inten, c, col).tobytes() == R.make_light(True, inten, c, col).tobytes()
        k = rng.normal(size=20)
        assert S.make_object(k, 0.25, col).tobytes() == R.make_object(k, 0.25, col).tobytes()
    assert np.array_equal(S.surface_clebsch(), R.make_surface("clebsch"))
    assert np.array_equal(S.surface_cayley(), R.make_surface("cayley"))
    with pytest.raises(RuntimeError, match="Negative value for sphere radius"):
        R.make_surface("sphere", [0, 0, 0], s=-1.0)
    with pytest.raises(RuntimeError, match="Invalid color"):
        R.make_object(np.zeros(20), 0.0, [1.5, 0, 0])


def test_clebsch_quirk_and_cayley(pkg):
    from cuda_ray_tracer_b200 import scene as S
    k = S.surface_clebsch()
    assert k[S.COEF_INDEX["x3"]] == 81.0 and k[S.COEF_INDEX["y3"]] == 81.0
    assert k[S.COEF_INDEX["z3"]] == 0.0          # surface.cpp:44 never sets z3
    assert k[S.COEF_INDEX["c"]] == 1.0 and k[S.COEF_INDEX["xyz"]] == 54.0
    c = S.surface_cayley()
    assert np.count_nonzero(c) == 9 and c[S.COEF_INDEX["xy"]] == 2.0 and c[S.COEF_INDEX["x2y"]] == -5.0


@pytest.mark.parametrize("name", SCENES)
def test_bundled_scenes_load(pkg, scenes_dir, name):
    sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml"))
    assert sc.px_width > 0 and sc.px_height > 0 and len(sc.objects) >= 1 and len(sc.lights) >= 1
    assert sc.max_reflections == 5                                  # scene.cpp:6 default, no bundled scene sets it
    ref_dir = "/root/reference/scenes"
    if os.path.isdir(ref_dir):                                      # fixtures carry exactly the reference's values
        rs = pkg.load_from_file(os.path.join(ref_dir, name + ".yml"))
        assert rs.objects.tobytes() == sc.objects.tobytes() and rs.lights.tobytes() == sc.lights.tobytes()
        assert (rs.px_width, rs.px_height, rs.vertical_fov) == (sc.px_width, sc.px_height, sc.vertical_fov)
        assert np.array_equal(rs.bg_color, sc.bg_color)


def test_defaults_and_validation(pkg, tmp_path):
    from cuda_ray_tracer_b200 import scene as S
    base = {"width": 4, "height": 2, "fov": 90, "objects": [], "light_sources": []}
    sc = S.scene_from_dict(dict(base))
    assert sc.max_reflections == 5 and sc.bg_color.tolist() == [1.0, 1.0, 1.0]   # scene.cpp:6-7 (white, not black)
    assert sc.vertical_fov == 90 * 0.01745329251994329576923690768489
    assert sc.aspect_ratio() == 2.0
    d = dict(base, objects=[{"type": "sphere", "color": [1, 1, 1]}], light_sources=[{"type": "spherical", "position": [0, 1, 0]}])
    sc = S.scene_from_dict(d)
    assert np.array_equal(sc.objects[0]["coef"], S.surface_sphere([0, 0, 0], 1.0))  # defaults: centre 0, radius 1
    assert sc.objects[0]["reflection_ratio"] == 0.0 and sc.lights[0]["light_color"].tolist() == [1, 1, 1]
    for bad, msg in [
        ({k: v for k, v in base.items() if k != "width"}, "Value 'width' undefined"),
        (dict(base, objects={"a": 1}), "must be a sequence"),
        (dict(base, objects=[{"type": "torus", "color": [0, 0, 0]}]), "Unknown surface type: 'torus'"),
        (dict(base, objects=[{"type": "sphere"}]), "Value 'color' undefined"),
        (dict(base, objects=[{"type": "sphere", "color": [0, 0, 2]}]), "Invalid color"),
        (dict(base, objects=[{"type": "sphere", "color": [0, 0, 0], "radius": -1}]), "Negative value for sphere radius"),
        (dict(base, objects=[{"type": "sphere", "color": [0, 0, 0], "reflection_ratio": -0.1}]), "Negative value for object reflection ratio"),
        (dict(base, objects=[{"type": "polynomial", "color": [0, 0, 0]}]), "Value 'coefficients' undefined"),
        (dict(base, light_sources=[{"type": "spot"}]), "must be 'spherical' or 'directional'"),
        (dict(base, light_sources=[{"type": "directional"}]), "Value 'direction' undefined"),
        (dict(base, light_sources=[{"type": "spherical", "position": [0, 0, 0], "intensity": -2}]), "Negative value for light intensity"),
        (dict(base, bg_color=[0, -1, 0]), "Invalid color"),
    ]:
        with pytest.raises(S.SceneException, match=msg):
            S.scene_from_dict(bad)
    with pytest.raises(S.SceneException, match="Cannot read the file"):
        S.load_from_file(str(tmp_path / "nope.yml"))
    (tmp_path / "bad.yml").write_text("width: [1, 2\nheight")
    with pytest.raises(S.SceneException, match="YAML parser error"):
        S.load_from_file(str(tmp_path / "bad.yml"))
    # reflection_ratio above 1 is accepted: only < 0 is rejected (scene.cpp:12)
    S.scene_from_dict(dict(base, objects=[{"type": "plane", "color": [0, 0, 0], "reflection_ratio": 1.5}]))


def test_camera_matrix(pkg):
    from cuda_ray_tracer_b200 import camera
    m = camera.camera_matrix().reshape(4, 4).T                       # back to [row, col]
    assert np.allclose(m, np.eye(4), atol=1e-15)                     # initial pose = identity to ~6e-17
    pos = (3.0, 2.0, -5.0)
    m = camera.camera_matrix(pos, 80.0, -10.0).reshape(4, 4).T
    assert np.allclose(m[:3, 3], pos)                                # origin = M*(0,0,0,1) = position
    assert np.allclose(m[:3, 2], camera.direction(80.0, -10.0))      # camera +z looks along `direction`
    assert np.allclose(m[:3, :3].T @ m[:3, :3], np.eye(3), atol=1e-14)
