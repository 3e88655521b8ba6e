"""The f32 "preview" mode (rtx_scene_render_preview_f32): the arithmetic of the reference's wgpu backend
(raytracing/gpu/base_shader.wgsl + the shape snippets) as a CUDA kernel, against its CPU restatement
oracle/wgsl_oracle.cpp.  The gate is bit-exact f32 (the kernel is built -fmad=false, the oracle -ffp-contract=off;
sin/cos are the f64 libm values rounded to f32 on both sides).  PARITY UNPINNED: the shader itself cannot be executed
on this box (no wgpu), so the oracle is one reading of it with nothing reference-held behind it — the CPU tests below
pin what can be pinned by hand: the RNG of base_shader.wgsl:99-132 and the loop bounds."""
import numpy as np
import pytest


def _scene(rtx, spp=2, bounces=4):
    from rtx_b200 import scenes
    return scenes.sphere_demo(rays_per_pixel=spp, max_bounces=bounces)


def _mixed(rtx, spp=2, bounces=5):
    from rtx_b200 import Camera, Config, Material, Object, Plane, Scene, Sphere, Triangle
    sc = Scene(Config().with_rays_per_pixel(spp).with_max_bounces(bounces), Camera((0, 0, 1.5), (1, 0.1, -0.05), 1.3))
    sc.add_object(Object(Plane((0, 0, 0), (0, 0, 2.5)), Material((0.7, 0.7, 0.6), (0, 0, 0), 0.8)))   # raw normal != unit
    sc.add_object(Object(Triangle([(6, -3, 0), (6, 3, 0), (6, 0, 4)]), Material((0.9, 0.3, 0.3), (0.2, 0, 0), 0.3)))
    sc.add_object(Object(Triangle([(3, 2, 0.1), (5, 4, 0.1), (4, 3, 3)]), Material((-0.2, 0.8, 0.3), (0, 0, 0), 0.0)))  # negative base: max(., 0)
    sc.add_object(Object(Sphere((4, -1.5, 1), 1.0), Material((1, 1, 1), (0, 0, 0), 0.0)))
    sc.add_object(Object(Sphere((3, 0.5, 0.4), 0.4), Material.light((3, 2.5, 2))))
    sc.add_object(Object(Sphere((2, 0, 8), 3), Material.light((2, 2, 2))))
    sc.add_object(Object(Plane((12, 0, 0), (-1, 0, 0)), Material((0.5, 0.5, 0.9), (0.05, 0.05, 0.05), 1.0)))
    return sc


# ---- CPU: the oracle against hand-restated pieces of the shader ----------------------------------------------------
def _pcg_seed(ux, uy):
    """base_shader.wgsl:99-123 in numpy scalars."""
    f = np.float32
    M = 0xFFFFFFFF
    ai = int(f(min(max(f(ux) + f(0.5), f(0.0)), f(0.999999))) * f(4294967296.0))
    bi = int(f(min(max(f(uy) + f(0.5), f(0.0)), f(0.999999))) * f(4294967296.0))
    h = ai ^ ((bi * 0x85ebca6b) & M)
    h = ((h ^ (h >> 16)) * 0x85ebca6b) & M
    x = h ^ (h >> 13)
    x = (x + (x << 10)) & M
    x ^= x >> 6
    x = (x + (x << 3)) & M
    x ^= x >> 11
    x = (x + (x << 15)) & M
    return x


def _pcg_next(seed):
    M = 0xFFFFFFFF
    s = (seed * 747796405 + 2891336453) & M
    word = (((s >> ((s >> 28) + 4)) ^ s) * 277803737) & M
    return s, (word >> 22) ^ word


def test_oracle_loop_bounds_and_first_sample(rtx, oracle):
    """max_bounces = 0: trace_ray's loop body never runs (base_shader.wgsl:153) — black image, no ray.  One bounce on
    an emitter that fills the view: every pixel = its emission (:160) and one ray per sample."""
    from rtx_b200 import Camera, Config, Material, Object, Scene, Sphere
    sc = _scene(rtx, spp=3, bounces=0)
    r = oracle.render_preview_scene(sc, 40, 30)
    assert r["rays"] == 0 and not r["rgb"].any() and (r["ids"] == 0xFFFFFFFF).all()
    sc = Scene(Config().with_rays_per_pixel(5).with_max_bounces(1), Camera((0, 0, 0), (1, 0, 0), 1.0))
    sc.add_object(Object(Sphere((10, 0, 0), 9), Material.light((0.25, 0.5, 2.0))))
    r = oracle.render_preview_scene(sc, 16, 12)
    assert r["rays"] == 16 * 12 * 5 and (r["ids"] == 0).all()
    assert np.array_equal(r["rgb"], np.broadcast_to(np.float32([0.25, 0.5, 2.0]), (12, 16, 3)))


def test_oracle_rng_matches_hand_restatement(rtx, oracle):
    """uv -> initRngSeed -> random_int -> random_float (base_shader.wgsl:99-132) restated in numpy scalars, against the
    oracle's generator, for the uv of every pixel of a small frame and a few edge values."""
    f = np.float32
    w, h = 7, 5
    uvs = [(f(f(f(x) + f(0.5)) / f(w)) - f(0.5), f(0.5) - f(f(f(row) + f(0.5)) / f(h))) for row in range(h) for x in range(w)]
    uvs += [(f(-0.5), f(0.5)), (f(0.5), f(-0.5)), (f(0.4999999), f(0.0)), (f(0.7), f(-0.9))]  # the clamp of :104-105
    for ux, uy in uvs:
        seed, u, fl = oracle.wgsl_rng_probe(ux, uy, 6)
        s = _pcg_seed(ux, uy)
        assert seed == s, (ux, uy)
        for i in range(6):
            s, r = _pcg_next(s)
            assert int(u[i]) == r
            assert fl[i] == f(f(r) / f(4294967296.0)) and 0.0 <= fl[i] <= 1.0


def test_oracle_is_deterministic_and_threading_independent(rtx, oracle):
    sc = _mixed(rtx)
    a = oracle.render_preview_scene(sc, 64, 40, nthreads=1)
    b = oracle.render_preview_scene(sc, 64, 40, nthreads=7)
    assert a["rgb"].tobytes() == b["rgb"].tobytes() and np.array_equal(a["ids"], b["ids"]) and a["rays"] == b["rays"]
    assert np.isfinite(a["rgb"]).all() and a["rgb"].max() > 0
    assert set(np.unique(a["ids"]).tolist()) >= {0, 1, 3}


# ---- GPU: the kernel against the oracle, bit for bit ----------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("w,h,spp,bounces", [(320, 200, 4, 5), (333, 77, 1, 8), (17, 5, 3, 2), (1, 1, 2, 3), (64, 64, 2, 0)])
def test_preview_kernel_equals_wgsl_oracle_s1(rtx, oracle, gpu, w, h, spp, bounces):
    sc = _scene(rtx, spp, bounces)
    rgb, ids = sc.render_preview(w, h, want_ids=True)
    ref = oracle.render_preview_scene(sc, w, h)
    assert rgb.dtype == np.float32 and rgb.shape == (h, w, 3)
    assert np.array_equal(ids, ref["ids"])
    assert rgb.tobytes() == ref["rgb"].tobytes(), f"{(rgb != ref['rgb']).sum()} of {rgb.size} channels differ"
    assert sc.last_stats.rays == ref["rays"] and sc.last_stats.kernel_launches == 1


@pytest.mark.gpu
def test_preview_kernel_mixed_shapes_and_updates(rtx, oracle, gpu):
    """Planes (raw shading normal), Moeller-Trumbore triangles, negative base colours, mirrors; then an in-place
    update and an added object: the preview's copy of the scene follows the object list."""
    from rtx_b200 import Material, Object, Sphere
    sc = _mixed(rtx)
    for (w, h) in ((200, 120), (31, 9)):
        rgb, ids = sc.render_preview(w, h, want_ids=True)
        ref = oracle.render_preview_scene(sc, w, h)
        assert np.array_equal(ids, ref["ids"]) and rgb.tobytes() == ref["rgb"].tobytes()
        assert sc.last_stats.rays == ref["rays"]
    sc.add_object(Object(Sphere((2.5, -0.8, 0.3), 0.3), Material((0.2, 0.9, 0.9), (0, 0, 0), 0.5)))
    rgb2 = sc.render_preview(200, 120)
    ref2 = oracle.render_preview_scene(sc, 200, 120)
    assert rgb2.tobytes() == ref2["rgb"].tobytes() and rgb2.tobytes() != rgb.tobytes()
    # the f64 path of the same handle is untouched by the preview
    a = sc.render(64, 40, seed=2)
    assert float(np.abs(a - oracle.render_scene(sc, 64, 40, seed=2)["rgb"]).max()) <= 1e-4


@pytest.mark.gpu
def test_preview_kernel_large_scene_from_global_memory(rtx, oracle, gpu):
    """More objects than fit the shared-memory stage (96 KB / 144 B per object): the nine arrays are read from HBM."""
    from rtx_b200 import scenes
    sc = scenes.random_spheres(n=1500, rays_per_pixel=1, max_bounces=3)
    rgb, ids = sc.render_preview(96, 54, want_ids=True)
    ref = oracle.render_preview_scene(sc, 96, 54)
    assert np.array_equal(ids, ref["ids"]) and rgb.tobytes() == ref["rgb"].tobytes() and sc.last_stats.rays == ref["rays"]


@pytest.mark.gpu
def test_preview_device_variant(rtx, oracle, gpu):
    import torch
    dev = torch.device("cuda", gpu)
    sc = _scene(rtx, 2, 4)
    w, h = 128, 72
    out = torch.full((h, w, 3), float("nan"), dtype=torch.float32, device=dev)
    cnt = torch.zeros(8, dtype=torch.int64, device=dev)
    sc.render_preview_device(w, h, out.data_ptr(), None, cnt.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize(dev)
    ref = oracle.render_preview_scene(sc, w, h)
    assert out.cpu().numpy().tobytes() == ref["rgb"].tobytes() and int(cnt[0].item()) == ref["rays"]


@pytest.mark.gpu
def test_preview_kernel_reproduces_golden_fixture(rtx, gpu):
    """The committed fixture (tests/golden/, generated from wgsl_oracle.cpp by make_golden.py), not a live oracle run."""
    import os
    from rtx_b200 import scenes
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "preview_sphere_demo_96x64_2spp.npz"))
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=5)
    rgb, ids = sc.render_preview(96, 64, want_ids=True)
    assert rgb.tobytes() == g["rgb"].tobytes() and np.array_equal(ids, g["ids"]) and sc.last_stats.rays == int(g["rays"])
