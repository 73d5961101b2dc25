"""CPU tests of the host scene builder against the reference's constants."""
import numpy as np
import pytest

from gpuraytracer_b200 import Renderer, _capi

# idx: (hit group kind, type, reflectance, step_scale, aabb min, aabb max)  -- src/Renderer.cpp:222-248, 505-535
REFERENCE_PRIMS = [
    (1, 0, 0.0, 1.0, (5, -1, -7), (7, 2, -5)),
    (1, 1, 1.0, 1.0, (2, -1, -4), (5, 2, -1)),
    (2, 0, 1.0, 1.0, (-7, -1, -7), (-4, 2, -4)),
    (3, 0, 0.0, 1.0, (1, -1, -7), (3, 1, -5)),
    (3, 1, 0.0, 1.0, (-7, -1, 1), (-5, 1, 3)),
    (3, 2, 1.0, 1.0, (-4, -1.4, 2), (-1, 1.6, 5)),
    (3, 3, 0.0, 0.5, (-7, -1, -3), (-5, 1, -1)),
    (3, 4, 0.0, 1.0, (-3, -1, -7), (-1, 1, -5)),
    (3, 5, 0.0, 1.0, (-7, -1, 5), (-5, 2, 7)),
    (3, 6, 0.0, 0.8, (1, -1, 1), (7, 5, 7)),
]


@pytest.fixture(scope="module")
def ref_scene():
    return Renderer(scene_id=0, device=None).scene()


def test_default_window_size(ref_scene):
    assert (ref_scene.width, ref_scene.height) == (1280, 720)  # src/main.cpp:14


def test_reference_primitives(ref_scene):
    s = ref_scene
    assert len(s.aabbs) == 10 and len(s.per_frame) == 10 and len(s.hit_groups) == 22
    for i, (kind, typ, refl, step, lo, hi) in enumerate(REFERENCE_PRIMS):
        np.testing.assert_allclose(s.aabbs[i]["min"], lo, rtol=0, atol=1e-6)
        np.testing.assert_allclose(s.aabbs[i]["max"], hi, rtol=0, atol=1e-6)
        for ray_type in (0, 1):  # record index = rayType + 2 * geometryIndex + 2
            rec = s.hit_groups[2 + 2 * i + ray_type]
            assert rec["shader_id"][0] == kind and rec["shader_id"][1] == ray_type
            assert rec["instance_index"] == i and rec["primitive_type"] == typ
            assert rec["material_cb"]["reflectance_coefficient"] == np.float32(refl)
            assert rec["material_cb"]["step_scale"] == np.float32(step)
    plane = s.hit_groups[0]["material_cb"]
    assert plane["reflectance_coefficient"] == np.float32(0.25) and plane["specular_coefficient"] == np.float32(0.4)
    np.testing.assert_allclose(plane["albedo"], [0.9, 0.9, 0.9, 1.0], rtol=1e-7)
    assert s.hit_groups[0]["shader_id"][0] == 0 and s.hit_groups[1]["shader_id"][1] == 1


def test_plane_and_instances(ref_scene):
    s = ref_scene
    assert list(s.indices) == [3, 1, 0, 2, 1, 3]  # src/Renderer.cpp:546-548
    np.testing.assert_array_equal(s.vertices["normal"], np.tile([0, 1, 0], (4, 1)).astype(np.float32))
    tri, box = s.instances
    assert tri["acceleration_structure"] == 0 and box["acceleration_structure"] == 1
    assert tri["hit_group_contrib_and_flags"] & 0xFFFFFF == 0 and box["hit_group_contrib_and_flags"] & 0xFFFFFF == 2
    assert tri["instance_id_and_mask"] >> 24 == 1
    t = tri["transform"]
    np.testing.assert_allclose([t[0, 0], t[1, 1], t[2, 2]], [2798, 2, 2798])          # 700*2 + 699*2
    np.testing.assert_allclose([t[0, 3], t[1, 3], t[2, 3]], [-979.3, 0, -979.3], rtol=1e-6)
    np.testing.assert_allclose(box["transform"], [[1, 0, 0, 0], [0, 1, 0, 1], [0, 0, 1, 0]])


def test_instanced_scene_tlas_and_hit_groups():
    """Scene 7 (two-level instancing): the reference primitives under a TLAS with three AABB instances and
    two triangle instances; the first two TLAS entries and the first 22 records are the reference's."""
    s = Renderer(scene_id=7, device=None).scene()
    ref = Renderer(scene_id=0, device=None).scene()
    assert len(s.instances) == 5 and len(s.aabbs) == 10
    assert [int(i["acceleration_structure"]) for i in s.instances] == [0, 1, 1, 1, 0]
    assert [int(i["hit_group_contrib_and_flags"]) & 0xFFFFFF for i in s.instances] == [0, 2, 2, 22, 42]
    np.testing.assert_array_equal(s.instances[:2]["transform"], ref.instances["transform"])
    assert len(s.hit_groups) == 2 + 20 + 20 + 2
    assert s.hit_groups[:22].tobytes() == ref.hit_groups.tobytes()
    a, b = s.hit_groups[2:22], s.hit_groups[22:42]
    np.testing.assert_array_equal(a["shader_id"], b["shader_id"])
    np.testing.assert_array_equal(a["instance_index"], b["instance_index"])
    # the two 2x2x2 SDF primitives swap their intersection type in the second block
    swapped = a["primitive_type"] != b["primitive_type"]
    assert sorted(np.flatnonzero(swapped) // 2) == [3, 3, 4, 4]
    # rotation + uniform scale 0.7 of the second AABB instance: rows of the 3x3 part have length 0.7
    t = s.instances[2]["transform"][:, :3]
    np.testing.assert_allclose(np.linalg.norm(t, axis=1), 0.7, rtol=1e-6)


def test_camera_and_light(ref_scene):
    cb = ref_scene.scene_cb[0]
    np.testing.assert_allclose(cb["camera_position"], [-12.0208153, 5.3, -12.0208153, 1.0], rtol=1e-6)
    np.testing.assert_allclose(cb["light_position"], [0, 18, -20, 0])
    np.testing.assert_allclose(cb["light_ambient_color"], [0.25, 0.25, 0.25, 1])
    np.testing.assert_allclose(cb["light_diffuse_color"], [0.6, 0.6, 0.6, 1])
    # projection_to_world maps NDC (0,0,0,1) onto the near plane on the eye -> at axis
    p = np.array([0, 0, 0, 1], np.float64) @ cb["projection_to_world"].astype(np.float64)
    p = p[:3] / p[3]
    eye = cb["camera_position"][:3].astype(np.float64)
    d = (p - eye) / np.linalg.norm(p - eye)
    np.testing.assert_allclose(d, -eye / np.linalg.norm(eye), atol=1e-4)
    assert abs(np.linalg.norm(p - eye) - 0.01) < 1e-4


def test_per_frame_transforms_are_inverse_pairs():
    r = Renderer(scene_id=0, device=None)
    for t in (0.0, 0.7, 3.0):
        r.on_update(t)
        s = r.scene()
        assert s.scene_cb[0]["elapsed_time"] == np.float32(t)
        for i in range(10):
            m = s.per_frame[i]["local_space_to_bottom_level_as"].astype(np.float64)
            inv = s.per_frame[i]["bottom_level_as_to_local_space"].astype(np.float64)
            np.testing.assert_allclose(m @ inv, np.eye(4), atol=2e-6)
            centre = 0.5 * (s.aabbs[i]["min"] + s.aabbs[i]["max"])
            np.testing.assert_allclose(m[3, :3], centre, atol=1e-6)  # translation row = AABB centre
    # rotating primitives (spheres, metaballs, twisted torus, cog) turn by -2t about y: XMMatrixRotationY
    r.on_update(0.25)
    m = r.scene().per_frame[1]["local_space_to_bottom_level_as"]
    c, s_ = np.cos(-0.5), np.sin(-0.5)
    np.testing.assert_allclose(m[:3, :3], 1.5 * np.array([[c, 0, -s_], [0, 1, 0], [s_, 0, c]]), atol=1e-6)
    assert np.allclose(r.scene().per_frame[0]["local_space_to_bottom_level_as"][:3, :3], np.diag([1, 1.5, 1]))


@pytest.mark.parametrize("scene_id,size,n_prims,depth,mb", [(1, (640, 480), 1, 2, 3), (2, (1920, 1080), 10, 3, 3), (3, (3840, 2160), 16, 3, 3),
                                                           (4, (1920, 1080), 9, 3, 5), (6, (3840, 2160), 16, 3, 3)])
def test_baseline_configs(scene_id, size, n_prims, depth, mb):
    s = Renderer(scene_id=scene_id, device=None).scene()
    assert (s.width, s.height) == size and len(s.aabbs) == n_prims
    assert s.config.max_recursion_depth == depth and s.config.metaballs_count == mb
    if scene_id == 1:
        assert all(h["material_cb"]["reflectance_coefficient"] == 0 for h in s.hit_groups)
    if scene_id == 3:
        assert (s.hit_groups[2::2]["primitive_type"] == 6).sum() >= 8  # FractalPyramid in >= 50 % of the cells
    if scene_id == 6:
        assert (s.config.sdf_max_steps, s.config.metaball_max_steps) == (2048, 512)


def test_config5_grid_is_parametric():
    s = Renderer(scene_id=5, grid_n=20, device=None).scene()
    assert (s.width, s.height) == (7680, 4320) and len(s.aabbs) == 400 and len(s.hit_groups) == 802
    kinds = s.hit_groups[2::2]["shader_id"][:, 0]
    assert set(kinds.tolist()) == {1, 2, 3}
    size = s.aabbs["max"] - s.aabbs["min"]
    assert np.all(size == 2.0)
    # cells do not overlap: stride 4, width 2
    xs = np.unique(s.aabbs["min"][:, 0])
    assert np.all(np.diff(xs) == 4.0)


def test_resize_updates_camera_only():
    r = Renderer(scene_id=0, device=None)
    a = r.scene().scene_cb[0]["projection_to_world"].copy()
    r.on_size_changed(640, 360)
    s = r.scene()
    assert (s.width, s.height) == (640, 360)
    np.testing.assert_allclose(s.scene_cb[0]["projection_to_world"], a, rtol=1e-5)  # same 16:9 aspect
    r.on_size_changed(640, 480)
    assert not np.allclose(r.scene().scene_cb[0]["projection_to_world"], a)


def test_camera_and_light_orbit():
    r = Renderer(scene_id=0, device=None)
    e0 = r.scene().scene_cb[0]["camera_position"].copy()
    r.on_update(0.0, camera_orbit_deg=90.0, light_orbit_deg=-45.0)
    cb = r.scene().scene_cb[0]
    assert np.isclose(np.linalg.norm(cb["camera_position"][:3]), np.linalg.norm(e0[:3]), rtol=1e-6)
    assert np.isclose(cb["camera_position"][1], e0[1])
    assert not np.allclose(cb["camera_position"], e0)
    assert np.isclose(np.linalg.norm(cb["light_position"][:3]), np.hypot(18, 20), rtol=1e-6)
