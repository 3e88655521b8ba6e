"""GPU: in-place object updates (rtx_scene_update_*): only the touched records are re-uploaded
and the LBVH is refitted — same hierarchy, new boxes — instead of rebuilt.  Gate: the image after an update
equals, byte for byte, the image of a scene built from scratch with the final data; a change of tree membership
falls back to a rebuild."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _fresh_like(rtx, sc):
    """A new scene holding sc's current host-side data."""
    from rtx_b200 import Scene
    types, shapes, mats = sc.objects_flat()
    new = Scene(sc.config, sc.camera, sc.device)
    ncol = {0: 4, 1: 6, 2: 9}
    for k, s, m in zip(types.tolist(), shapes, mats):
        new._append(int(k), s[None, :ncol[int(k)]].copy(), m[None, :].copy())
    return new


def _same(a, b):
    return a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])


def _equals_oracle(oracle, sc, img, w, h, seed):
    """The updated scene's image against the CPU oracle's render of the scene's CURRENT host-side data (the object list
    a reference user holds after mutating the shapes behind their Arc<Mutex<..>>, object.rs:12): primary hit ids and
    ray count equal, every radiance byte equal."""
    ref = oracle.render_scene(sc, w, h, seed=seed)
    assert np.array_equal(img[1], ref["ids"]), f"{(img[1] != ref['ids']).sum()} primary hit ids differ from the oracle"
    assert sc.last_stats.rays == ref["rays"], (sc.last_stats.rays, ref["rays"])
    assert img[0].tobytes() == ref["rgb"].tobytes(), float(np.abs(img[0] - ref["rgb"]).max())


def test_sphere_updates_refit_the_lbvh(rtx, oracle, gpu):
    from rtx_b200 import scenes
    LB, BR = rtx.RTX_ACCEL_LBVH, rtx.RTX_ACCEL_BRUTE
    sc = scenes.random_spheres(n=2000, rays_per_pixel=2, max_bounces=4)
    w, h = 240, 135
    before = sc.render(w, h, seed=3, flags=LB, want_ids=True)
    assert (sc.accel_builds, sc.accel_refits) == (1, 0)
    types, shapes, mats = sc.objects_flat()
    rng = np.random.default_rng(1)
    idx = rng.choice(2000, size=300, replace=False)
    new_shapes = shapes[idx, :4].copy()
    new_shapes[:, :3] += rng.normal(size=(300, 3)) * 0.7   # move
    new_shapes[:, 3] *= rng.uniform(0.5, 1.5, size=300)    # resize
    new_mats = mats[idx].copy()
    new_mats[:, :3] = rng.uniform(0.1, 0.9, size=(300, 3))
    sc.update_objects(idx, new_shapes, new_mats)
    after = sc.render(w, h, seed=3, flags=LB, want_ids=True)
    assert (sc.accel_builds, sc.accel_refits) == (1, 1), "the update must refit, not rebuild"
    assert not _same(before, after)
    _equals_oracle(oracle, sc, after, w, h, 3)
    fresh = _fresh_like(rtx, sc)
    assert _same(after, fresh.render(w, h, seed=3, flags=LB, want_ids=True))
    assert _same(after, fresh.render(w, h, seed=3, flags=BR, want_ids=True))
    # shapes only (materials kept), a second refit, and then a forced rebuild: same image each time
    sc.update_objects(idx[:10], new_shapes[:10] + 0.25)
    again = sc.render(w, h, seed=3, flags=LB, want_ids=True)
    assert (sc.accel_builds, sc.accel_refits) == (1, 2)
    _equals_oracle(oracle, sc, again, w, h, 3)
    assert _same(again, _fresh_like(rtx, sc).render(w, h, seed=3, flags=BR, want_ids=True))
    sc.rebuild_accel()
    assert _same(again, sc.render(w, h, seed=3, flags=LB, want_ids=True)) and sc.accel_builds == 2


def test_membership_change_falls_back_to_rebuild(rtx, oracle, gpu):
    from rtx_b200 import scenes
    LB = rtx.RTX_ACCEL_LBVH
    sc = scenes.random_spheres(n=500, rays_per_pixel=1, max_bounces=3)
    sc.render(160, 90, seed=1, flags=LB)
    types, shapes, mats = sc.objects_flat()
    bad = shapes[7, :4].copy()
    bad[0] = np.nan  # a sphere that can never be hit any more: it leaves the tree
    sc.update_objects([7], bad[None, :])
    img = sc.render(160, 90, seed=1, flags=LB, want_ids=True)
    assert sc.accel_builds == 2
    _equals_oracle(oracle, sc, img, 160, 90, 1)
    assert _same(img, _fresh_like(rtx, sc).render(160, 90, seed=1, flags=rtx.RTX_ACCEL_BRUTE, want_ids=True))


def test_triangle_updates(rtx, oracle, gpu):
    from rtx_b200 import scenes
    LB, BR = rtx.RTX_ACCEL_LBVH, rtx.RTX_ACCEL_BRUTE
    sc = scenes.terrain(grid=25, rays_per_pixel=1, max_bounces=4)
    w, h = 160, 90
    sc.render(w, h, seed=2, flags=LB)
    types, shapes, mats = sc.objects_flat()
    tri = np.flatnonzero(types == 2)
    rng = np.random.default_rng(5)
    idx = rng.choice(tri, size=80, replace=False)
    moved = shapes[idx].copy()
    moved[:, 2::3] += rng.normal(size=(80, 3)) * 0.05  # z of each vertex: same pivots, new planes and boxes
    sc.update_objects(idx, moved)
    img = sc.render(w, h, seed=2, flags=LB, want_ids=True)
    assert (sc.accel_builds, sc.accel_refits) == (1, 1)
    _equals_oracle(oracle, sc, img, w, h, 2)
    fresh = _fresh_like(rtx, sc)
    assert _same(img, fresh.render(w, h, seed=2, flags=LB, want_ids=True))
    assert _same(img, fresh.render(w, h, seed=2, flags=BR, want_ids=True))
    # a triangle collapsing to a segment changes its elimination code ("can't handle LGS"): rebuild
    deg = shapes[tri[3]].copy()
    deg[3:6] = deg[0:3]
    deg[6:9] = deg[0:3] + np.array([0.0, 1.0, 0.0])
    sc.update_objects([tri[3]], deg[None, :])
    img2 = sc.render(w, h, seed=2, flags=LB, want_ids=True)
    assert sc.accel_builds == 2
    _equals_oracle(oracle, sc, img2, w, h, 2)
    assert _same(img2, _fresh_like(rtx, sc).render(w, h, seed=2, flags=BR, want_ids=True))


def test_small_scene_partial_upload_and_planes(rtx, oracle, gpu):
    """S1 (pool kernel, scanned from shared memory): updates travel as partial uploads of the scene blob."""
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=8)
    a = sc.render(200, 120, seed=1, want_ids=True)
    types, shapes, mats = sc.objects_flat()
    sc.update_objects([1], np.array([[6.0, 0.3, 1.2, 0.8]]), np.array([[0.1, 0.8, 0.3, 0, 0, 0, 0.4]]))
    sc.update_objects([0], np.array([[0.0, 0.0, -0.2, 0.0, 0.1, 1.0]]))  # the ground plane, tilted
    b = sc.render(200, 120, seed=1, want_ids=True)
    assert not _same(a, b)
    _equals_oracle(oracle, sc, b, 200, 120, 1)
    assert _same(b, _fresh_like(rtx, sc).render(200, 120, seed=1, want_ids=True))


def test_update_errors(rtx, gpu):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=1)
    sc.render(32, 32, seed=1)
    with pytest.raises(rtx.RtxError):
        sc.update_objects([99], np.zeros((1, 4)))
    with pytest.raises(rtx.RtxError):  # object 0 is a plane
        sc.update_objects([0], np.zeros((1, 4)))
    with pytest.raises(rtx.RtxError):  # mixed kinds
        sc.update_objects([0, 1], np.zeros((2, 4)))
