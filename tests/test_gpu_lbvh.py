"""GPU: the LBVH (K4 build + K5 stackless traversal) must return exactly what the O(N) scan of
scene.rs:243-251 returns — same distance bits, same object index (lowest index on ties), including the
reference's phantom triangle hits.  Gate: LBVH == brute-force kernel, bit for bit; brute force itself is
gated against the oracle in test_gpu_functions.py / test_gpu_render.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rays(rng, n, spread):
    pos = rng.uniform(-spread, spread, (n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([pos, d], axis=1)


def _awkward_rays(rng, n, spread):
    """Rays the f32 slab test has to survive: axis-parallel and nearly axis-parallel directions (1/d not
    representable or huge), origins far outside the scene, a NaN direction."""
    r = _rays(rng, n, spread)
    q = n // 8
    for k in range(3):
        r[k * q:(k + 1) * q, 3 + k] = 0.0                      # one zero component
    r[3 * q:4 * q, 3:] = np.eye(3)[rng.integers(0, 3, q)] * rng.choice([-1.0, 1.0], (q, 1))  # axis-parallel
    r[4 * q:5 * q, 3] = rng.uniform(-1e-12, 1e-12, q)          # almost zero component
    r[5 * q:6 * q, 4] = rng.choice([-1.0, 1.0], q) * 1e-300    # 1/d overflows f32
    r[6 * q:7 * q, :3] *= 1.0e6                                # far origins
    d = r[:, 3:]
    r[:, 3:] = d / np.linalg.norm(d, axis=1, keepdims=True)
    r[-1, 3:] = np.nan
    r[-2, :3] = np.nan
    return r


def _same(sc, rtx, rays):
    tb, ib = sc.debug_closest_object(rays, flags=rtx.RTX_ACCEL_BRUTE)
    tl, il = sc.debug_closest_object(rays, flags=rtx.RTX_ACCEL_LBVH)
    assert np.array_equal(ib, il), f"{(ib != il).sum()} of {len(ib)} hit ids differ"
    assert tb.tobytes() == tl.tobytes()
    return ib


def _visible_triangles(rng, n, spread=20.0, size=2.0):
    out = []
    while len(out) < n:
        c = rng.uniform(-spread, spread, 3)
        v = c + rng.uniform(-size, size, (3, 3))
        nrm = np.cross(v[1] - v[0], v[2] - v[0])
        ln = np.linalg.norm(nrm)
        if ln < 1e-3:
            continue
        if (nrm / ln) @ v[0] < 0:
            v = v[[0, 2, 1]]
        out.append(v.ravel())
    return np.array(out)


def _mats(n):
    return np.tile([0.6, 0.6, 0.6, 0, 0, 0, 1.0], (n, 1))


def test_spheres_lbvh_equals_scan(rtx, gpu):
    rng = np.random.default_rng(21)
    n = 3000
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    centers = rng.uniform(-30, 30, (n, 3))
    radii = rng.uniform(0.05, 2.0, n)
    centers[100:110] = centers[0:10]  # exact duplicates → distance ties → lowest index must win
    radii[100:110] = radii[0:10]
    sc.add_spheres(centers, radii, _mats(n))
    rays = _rays(rng, 60000, 35.0)
    # rays starting exactly on surfaces and at centres
    k = 2000
    nrm = rng.normal(size=(k, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    rays[:k, :3] = centers[:k] + nrm * radii[:k, None]
    rays[k:k + 500, :3] = centers[:500]
    ids = _same(sc, rtx, rays)
    assert (ids != rtx.MISS_ID).mean() > 0.3
    dup = set(range(100, 110))
    assert not (dup & set(ids.tolist()))  # the higher-index duplicate never wins


def test_triangles_lbvh_equals_scan_including_phantoms(rtx, oracle, gpu):
    rng = np.random.default_rng(22)
    tris = _visible_triangles(rng, 1500)
    # other pivot classes: r.x == 0 (dropped axis differs) and vertical walls
    special = np.array([
        [0, 0, -25, 0, 4, -25, 4, 0, -25.0],
        [25, 0, 0, 25, 0, 4, 25, -4, 0.0],
        [2, 0, 13, 4, 0, 13, 2, 0, 17.0],
        [0, 0, 5, 0, 0, 5, 0, 4, 5.0],       # degenerate
    ])
    tris = np.concatenate([tris, special])
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_triangles(tris, _mats(len(tris)))
    rays = _rays(rng, 60000, 25.0)
    # bounce-like rays: origins ON triangles (self-hit noise |t| ~ 1e-16 must match bit for bit)
    k = 20000
    t = tris[rng.integers(0, 1500, k)].reshape(k, 3, 3)
    a = rng.uniform(0, 1, (k, 1))
    b = rng.uniform(0, 1, (k, 1)) * (1 - a)
    rays[:k, :3] = t[:, 0] + a * (t[:, 1] - t[:, 0]) + b * (t[:, 2] - t[:, 0])
    ids = _same(sc, rtx, rays)
    assert (ids != rtx.MISS_ID).mean() > 0.3
    # and the scan agrees with the oracle on a subset (ties the LBVH to the reference semantics)
    objs = oracle.make_objects(*sc.objects_flat())
    t_ref, i_ref = oracle.closest_object(objs, rays[:3000])
    t_l, i_l = sc.debug_closest_object(rays[:3000], flags=rtx.RTX_ACCEL_LBVH)
    assert np.array_equal(i_ref, i_l) and t_ref.tobytes() == t_l.tobytes()
    # a pure phantom configuration: everything is behind the rays
    ph = _rays(rng, 20000, 10.0)
    ph[:, 2] = 60.0
    ph[:, 5] = np.abs(ph[:, 5])
    ids = _same(sc, rtx, ph)
    assert (ids != rtx.MISS_ID).any()  # phantom hits exist and the prism trees find them


def test_prism_gradient_bound_stress(rtx, gpu):
    """The prism trees prune with per-child bounds of the triangles' plane gradients (rtx_lbvh.cuh,
    prism_gradient_test).  Configurations that stress it: planes almost parallel to the dropped axis (huge gradients),
    exactly parallel ones (no bound), slivers, a rough height field with steep cliffs, coordinates far from the origin,
    and phantom-heavy ray sets (origins below / behind / on the surface, grazing directions).  Ids and distance bits
    must equal the scan's."""
    rng = np.random.default_rng(77)
    # (1) a rough height field with cliffs: slopes from 0 to ~40
    g = 48
    xs = np.linspace(-12.0, 12.0, g)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    Z = 5.0 + 2.0 * np.sin(0.9 * X) * np.cos(0.7 * Y) + rng.uniform(-0.4, 0.4, (g, g)) + 6.0 * (X > 3.0)
    P = np.stack([X, Y, Z], axis=-1)
    p00, p10, p01, p11 = P[:-1, :-1], P[:-1, 1:], P[1:, :-1], P[1:, 1:]
    field = np.concatenate([np.stack([p00, p10, p11], axis=2).reshape(-1, 9), np.stack([p00, p11, p01], axis=2).reshape(-1, 9)])
    # (2) a soup with near-vertical and exactly axis-parallel planes and slivers, some far away
    soup = _visible_triangles(rng, 600, spread=15.0, size=3.0)
    walls = []
    for k in range(120):
        c = rng.uniform(-10, 10, 3)
        ax = k % 3
        e1, e2 = np.zeros(3), np.zeros(3)
        e1[(ax + 1) % 3] = rng.uniform(0.5, 3)
        e2[(ax + 2) % 3] = rng.uniform(0.5, 3)
        e2[ax] = 0.0 if k % 2 else rng.uniform(-1e-9, 1e-9)  # exactly / almost parallel to an axis
        walls.append(np.concatenate([c, c + e1, c + e2]))
    slivers = [np.concatenate([c, c + [3, 0, 1e-7], c + [6, 1e-7, 0]]) for c in rng.uniform(-8, 8, (40, 3))]
    far = _visible_triangles(rng, 100, spread=5.0, size=1.0) + np.tile([4000.0, -2500.0, 900.0], 3)
    tris = np.concatenate([field, soup, np.array(walls), np.array(slivers), far])
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_triangles(tris, _mats(len(tris)))
    n = 160000
    rays = _rays(rng, n, 16.0)
    q = n // 8
    # origins ON triangles, directions grazing (small component along the triangle's normal)
    t = tris[rng.integers(0, len(field), q)].reshape(q, 3, 3)
    a = rng.uniform(0, 1, (q, 1))
    b = rng.uniform(0, 1, (q, 1)) * (1 - a)
    rays[:q, :3] = t[:, 0] + a * (t[:, 1] - t[:, 0]) + b * (t[:, 2] - t[:, 0])
    graze = rng.normal(size=(q, 3))
    graze[:, 2] = rng.uniform(0.0, 0.3, q)
    rays[:q, 3:] = graze / np.linalg.norm(graze, axis=1, keepdims=True)
    # origins below the field looking up, above looking up, beside looking away: everything behind or ahead
    rays[q:2 * q, 2] = -3.0
    rays[q:2 * q, 5] = np.abs(rays[q:2 * q, 5])
    rays[2 * q:3 * q, 2] = 30.0
    rays[2 * q:3 * q, 5] = np.abs(rays[2 * q:3 * q, 5])
    rays[3 * q:4 * q, 0] = -40.0
    rays[3 * q:4 * q, 3] = -np.abs(rays[3 * q:4 * q, 3])
    # near the far cluster
    rays[4 * q:5 * q, :3] = rng.uniform(-8, 8, (q, 3)) + [4000.0, -2500.0, 900.0]
    d = rays[:, 3:]
    rays[:, 3:] = d / np.linalg.norm(d, axis=1, keepdims=True)
    ids = _same(sc, rtx, rays)
    assert (ids != rtx.MISS_ID).mean() > 0.2
    # phantom hits are among them: a hit whose plane lies behind the origin
    tb, ib = sc.debug_closest_object(rays[q:3 * q], flags=rtx.RTX_ACCEL_BRUTE)
    assert (ib != rtx.MISS_ID).any()


def test_mixed_scene_with_planes_and_invalid_prims(rtx, gpu):
    rng = np.random.default_rng(23)
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_planes([[0, 0, -12]], [[0, 0, 1]], _mats(1))
    sc.add_spheres(rng.uniform(-10, 10, (200, 3)), rng.uniform(0.1, 1.5, 200), _mats(200))
    sc.add_triangles(_visible_triangles(rng, 200, spread=10.0), _mats(200))
    sc.add_spheres([[np.nan, 0, 0], [0, 0, 0], [1, 2, 3]], [1.0, np.inf, np.nan], _mats(3))  # can never be hit
    sc.add_planes([[0, 0, 14]], [[0, 0, -1]], _mats(1))
    sc.add_spheres(rng.uniform(-10, 10, (100, 3)), rng.uniform(0.1, 1.5, 100), _mats(100))
    ids = _same(sc, rtx, _rays(rng, 50000, 11.0))
    seen = set(ids.tolist())
    assert 0 in seen and 404 in seen and not ({401, 402, 403} & seen)


def test_tiny_trees(rtx, gpu):
    rng = np.random.default_rng(24)
    for n in (1, 2, 3):
        sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
        sc.add_spheres(rng.uniform(-3, 3, (n, 3)), rng.uniform(0.5, 1.5, n), _mats(n))
        _same(sc, rtx, _rays(rng, 5000, 5.0))
        _same(sc, rtx, _awkward_rays(rng, 4000, 5.0))
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_triangles(_visible_triangles(rng, 1, spread=3.0), _mats(1))  # one-primitive trees (3-D and prism)
    _same(sc, rtx, _rays(rng, 5000, 5.0))
    _same(sc, rtx, _awkward_rays(rng, 4000, 5.0))


def test_awkward_rays(rtx, gpu):
    rng = np.random.default_rng(25)
    sc = rtx.Scene(rtx.Config(), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_spheres(rng.uniform(-10, 10, (400, 3)), rng.uniform(0.1, 1.5, 400), _mats(400))
    tris = _visible_triangles(rng, 400, spread=10.0)
    tris[:50, [2, 5, 8]] = np.round(tris[:50, [2]])     # axis-aligned (flat boxes), integer heights
    sc.add_triangles(tris, _mats(400))
    rays = _awkward_rays(rng, 40000, 11.0)
    rays[:2000, 2] = np.round(rays[:2000, 2])           # origins in the planes of the flat triangles
    ids = _same(sc, rtx, rays)
    assert (ids != rtx.MISS_ID).mean() > 0.2


@pytest.mark.parametrize("which", ["spheres", "terrain"])
def test_whole_image_lbvh_equals_scan(rtx, gpu, which):
    from rtx_b200 import scenes
    if which == "spheres":
        sc, (w, h) = scenes.random_spheres(n=2000, rays_per_pixel=2, max_bounces=8), (320, 180)
    else:
        sc, (w, h) = scenes.terrain(grid=64, rays_per_pixel=2, max_bounces=4), (256, 144)
    a, ia = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_BRUTE, want_ids=True)
    sa = (sc.last_stats.rays, sc.last_stats.hits, sc.last_stats.accel_used)
    b, ib = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_LBVH, want_ids=True)
    sb = (sc.last_stats.rays, sc.last_stats.hits, sc.last_stats.accel_used)
    assert a.tobytes() == b.tobytes() and np.array_equal(ia, ib)
    assert sa[:2] == sb[:2] and sa[2] == rtx.RTX_ACCEL_BRUTE and sb[2] == rtx.RTX_ACCEL_LBVH
    st = sc.last_stats
    assert st.nodes_visited > 0 and st.prims_tested < st.rays * sc.object_count / 10
    print(f"{which}: {st.rays} rays, {st.nodes_visited / st.rays:.1f} nodes/ray, {st.prims_tested / st.rays:.1f} prims/ray")
    c = sc.render(w, h, seed=1)  # AUTO picks the LBVH here
    assert sc.last_stats.accel_used == rtx.RTX_ACCEL_LBVH and c.tobytes() == b.tobytes()


def test_s2_10k_spheres_against_oracle(rtx, oracle, gpu):
    """Config C3's scene at a size the oracle's O(N) scan finishes in seconds."""
    from rtx_b200 import scenes
    sc = scenes.random_spheres(n=10_000, rays_per_pixel=1, max_bounces=8)
    w, h = 240, 135
    rgb, ids = sc.render(w, h, seed=1, want_ids=True)
    assert sc.last_stats.accel_used == rtx.RTX_ACCEL_LBVH
    ref = oracle.render_scene(sc, w, h, seed=1)
    assert np.array_equal(ids, ref["ids"]) and sc.last_stats.rays == ref["rays"]
    assert np.abs(rgb - ref["rgb"]).max() <= 1e-4


def test_s3_million_triangles_lbvh_vs_bruteforce_ids(rtx, gpu):
    """Config C4: 1 002 528-triangle terrain, primary + 4 bounces; hit ids of the LBVH render checked against the
    brute-force kernel on a 240x135 image (the full-depth brute force is 1e6 tests per ray)."""
    from rtx_b200 import scenes
    sc = scenes.terrain(grid=709, rays_per_pixel=1, max_bounces=4)
    assert sc.object_count == 1_002_529
    w, h = 240, 135
    b, ib = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_LBVH, want_ids=True)
    sb = sc.last_stats
    rays_l, ms_l, npr = sb.rays, sb.device_ms, sb.prims_tested / sb.rays
    a, ia = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_BRUTE, want_ids=True)
    assert np.array_equal(ia, ib) and a.tobytes() == b.tobytes() and sc.last_stats.rays == rays_l
    print(f"S3: {rays_l} rays; LBVH {ms_l:.1f} ms ({npr:.1f} prims/ray) vs scan {sc.last_stats.device_ms:.1f} ms")
    assert (ib != rtx.MISS_ID).mean() > 0.3


def test_c3_stated_size_subgrid_vs_oracle(rtx, oracle, gpu):
    """Config C3 at its STATED size — S2, 10 000 spheres, 3840x2160, 16 spp, max depth 8, LBVH — against the CPU
    oracle's O(N) scan on the pixel sub-grid SURVEY.md 8(d) names (every 16th pixel in x and y: 240x135 pixels x 16
    samples x <= 9 segments x 10^4 spheres, seconds on the box's host threads).  A pixel's RNG streams are keyed by
    its position in the FULL image ((y*W + x)*rpp + s), so the oracle's sub-grid pixels are the full render's pixels:
    ids and every radiance byte must be equal."""
    from rtx_b200 import scenes
    sc = scenes.random_spheres(n=10_000, rays_per_pixel=16, max_bounces=8)
    w, h, g = 3840, 2160, 16
    rgb, ids = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_LBVH, want_ids=True)
    st = sc.last_stats
    assert st.accel_used == rtx.RTX_ACCEL_LBVH and st.paths == w * h * 16
    ref = oracle.render_scene(sc, w, h, seed=1, xs=g, ys=g)
    assert ref["ids"].shape == (135, 240)
    assert np.array_equal(ids[::g, ::g], ref["ids"]), f"{(ids[::g, ::g] != ref['ids']).sum()} primary hit ids differ"
    sub = np.ascontiguousarray(rgb[::g, ::g])
    assert float(np.abs(sub - ref["rgb"]).max()) <= 1e-4  # the north-star gate ...
    assert sub.tobytes() == ref["rgb"].tobytes()          # ... and the bar this repo holds
    assert (ref["ids"] != rtx.MISS_ID).mean() > 0.2
    print(f"C3 stated size: {st.rays} rays, {st.device_ms:.1f} ms, {st.nodes_visited / st.rays:.1f} nodes/ray; "
          f"{ref['rays']} oracle rays on the {g}-grid")


def test_c4_stated_size_subgrid_vs_oracle(rtx, oracle, gpu):
    """Config C4 at its STATED size — S3, 1 002 528 triangles + light, 1920x1080, primary + 4 bounces, LBVH (3-D tree +
    prism trees) — against the CPU oracle's brute-force scan on a pixel sub-grid.  The as-written triangle test costs
    the oracle ~75 ns, i.e. 10^6 triangles x ~4.5 segments = 0.3 s per path per host thread, so the grid is every 32nd
    pixel (60x34 paths, under a minute on 16 threads) rather than SURVEY.md 8(d)'s every 8th (a quarter of an hour);
    test_s3_million_triangles_lbvh_vs_bruteforce_ids covers a 240x135 image against the device's own scan."""
    from rtx_b200 import scenes
    sc = scenes.terrain(grid=709, rays_per_pixel=1, max_bounces=4)
    w, h, g = 1920, 1080, 32
    rgb, ids = sc.render(w, h, seed=1, flags=rtx.RTX_ACCEL_LBVH, want_ids=True)
    st = sc.last_stats
    assert st.accel_used == rtx.RTX_ACCEL_LBVH and st.triangle_panics == 0
    ref = oracle.render_scene(sc, w, h, seed=1, xs=g, ys=g)
    assert np.array_equal(ids[::g, ::g], ref["ids"]), f"{(ids[::g, ::g] != ref['ids']).sum()} primary hit ids differ"
    sub = np.ascontiguousarray(rgb[::g, ::g])
    assert float(np.abs(sub - ref["rgb"]).max()) <= 1e-4
    assert sub.tobytes() == ref["rgb"].tobytes()
    assert (ref["ids"] != rtx.MISS_ID).mean() > 0.3 and ref["panics"] == 0
    print(f"C4 stated size: {st.rays} rays, {st.device_ms:.2f} ms, {st.nodes_visited / st.rays:.1f} nodes/ray, "
          f"{st.prims_tested / st.rays:.1f} prims/ray; {ref['rays']} oracle rays on the {g}-grid")
