"""GPU parity, whole image: the CUDA render path against the CPU oracle on the same scene and seed,
through the C ABI.  Gates (BASELINE.json north_star): primary-hit IDs bit-exact; linear radiance within
1e-4 max-abs per channel at 1 spp; PSNR >= 60 dB at N spp; ray counts equal."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_1SPP = 1e-4  # north_star: "within 1e-4 max abs error per channel at 1 spp"
PSNR_MIN = 60.0  # north_star: "PSNR >= 60 dB at N spp"


def psnr(a, b):
    peak = max(1.0, float(np.nanmax(b)))
    mse = float(np.mean((a - b) ** 2))
    return np.inf if mse == 0 else 10.0 * np.log10(peak * peak / mse)


def _compare(rtx, oracle, sc, w, h, seed, spp_tol=TOL_1SPP):
    rgb, ids = sc.render(w, h, seed=seed, want_ids=True)
    ref = oracle.render_scene(sc, w, h, seed=seed)
    assert np.array_equal(ids, ref["ids"]), f"{(ids != ref['ids']).sum()} primary hit ids differ"
    assert sc.last_stats.rays == ref["rays"], (sc.last_stats.rays, ref["rays"])
    err = float(np.abs(rgb - ref["rgb"]).max())
    assert err <= spp_tol, err
    return rgb, ids, ref, err


def test_sphere_demo_800x600_1spp(rtx, oracle, gpu):
    """Config C1: the demo scene at 800x600, 1 spp, max depth 8."""
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=8)
    rgb, ids, ref, err = _compare(rtx, oracle, sc, 800, 600, seed=1)
    st = sc.last_stats
    assert st.paths == 800 * 600 and st.pixels == 800 * 600 and st.kernel_launches == 1
    frac_exact = float(np.mean(rgb == ref["rgb"]))
    print(f"C1: max|err|={err:.3e}, bit-equal channels {frac_exact:.4f}, rays={st.rays}, {st.device_ms:.3f} ms")
    assert frac_exact > 0.9  # only paths that bounced (sin/cos) can differ at all


@pytest.mark.parametrize("w,h", [(256, 256), (333, 77), (17, 5), (1, 1), (512, 512), (1024, 1024), (1920, 1080),
                                 (2048, 2048), (3840, 2160), (4096, 4096), (7680, 4320)])
def test_primary_only_hit_ids_bit_exact(rtx, oracle, gpu, w, h):
    """Config C5, the whole sweep of BASELINE.json configs[4] (256^2 ... 7680x4320) plus ragged sizes: max_bounces = 0,
    1 spp — no libm on the device path, so radiance is bit-exact too (scene.rs:144-159, 213-222, 243-251)."""
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=0)
    rgb, ids, ref, err = _compare(rtx, oracle, sc, w, h, seed=1, spp_tol=0.0)
    assert rgb.tobytes() == ref["rgb"].tobytes()
    del rgb
    ids_only = sc.render_ids(w, h, seed=1)
    assert np.array_equal(ids_only, ref["ids"])


def test_psnr_at_16spp(rtx, oracle, gpu):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=16, max_bounces=8)
    rgb, ids = sc.render(480, 270, seed=1, want_ids=True)
    ref = oracle.render_scene(sc, 480, 270, seed=1)
    assert np.array_equal(ids, ref["ids"])
    assert sc.last_stats.rays == ref["rays"]
    p = psnr(rgb, ref["rgb"])
    print(f"16 spp PSNR = {p:.1f} dB, max|err| = {np.abs(rgb - ref['rgb']).max():.3e}")
    assert p >= PSNR_MIN


def test_deterministic_and_seed_sensitive(rtx, gpu):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=8)
    a = sc.render(320, 200, seed=7)
    b = sc.render(320, 200, seed=7)
    c = sc.render(320, 200, seed=8)
    assert a.tobytes() == b.tobytes() and a.tobytes() != c.tobytes()


def test_empty_scene_zero_spp_and_depth_limits(rtx, oracle, gpu):
    from rtx_b200 import scenes
    sc = rtx.Scene(rtx.Config().with_rays_per_pixel(2), rtx.Camera((0, 0, 0), (1, 0, 0), 1.0))
    rgb, ids = sc.render(33, 9, seed=0, want_ids=True)
    assert not rgb.any() and (ids == rtx.MISS_ID).all() and sc.last_stats.rays == 0  # scene.rs:224-226
    sc = scenes.sphere_demo(rays_per_pixel=0)
    rgb, ids = sc.render(20, 10, seed=0, want_ids=True)
    assert np.isnan(rgb).all() and (ids == rtx.MISS_ID).all() and sc.last_stats.rays == 0  # scene.rs:253-259
    for mb in (0, 1, 2, 10):
        sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=mb)
        _compare(rtx, oracle, sc, 96, 64, seed=3)


def _compare_triangle_scene(rtx, oracle, sc, w, h, seed):
    """Triangle scenes are the sensitive ones: a ray leaving a triangle re-tests that triangle with
    |t| = |n.(v0-pos) / (d.n)| (the abs() of triangle.rs:118), i.e. rounding noise ~1e-16 that is a self-hit
    whenever it is not exactly 0 — and whether it is depends on the last bit of the bounce direction.  With the
    device evaluating sin/cos by glibc's algorithm (rtx_sincos.cuh) every bit agrees: ids, ray count, radiance."""
    rgb, ids = sc.render(w, h, seed=seed, want_ids=True)
    ref = oracle.render_scene(sc, w, h, seed=seed)
    assert np.array_equal(ids, ref["ids"]), f"{(ids != ref['ids']).sum()} primary hit ids differ"
    differing = float(np.mean(np.any(rgb != ref["rgb"], axis=2)))
    print(f"triangle scene: pixels differing {differing:.2e}, rays {sc.last_stats.rays} vs {ref['rays']}")
    assert int(sc.last_stats.rays) == int(ref["rays"]), (sc.last_stats.rays, ref["rays"])
    assert rgb.tobytes() == ref["rgb"].tobytes(), differing
    return rgb, ids, ref


def test_triangles_planes_spheres_scene(rtx, oracle, gpu):
    from rtx_b200 import Camera, Config, Material, Object, Plane, Scene, Sphere, Triangle
    sc = Scene(Config().with_rays_per_pixel(2).with_max_bounces(4), Camera((0, 0, 1), (1, 0, 0.2), 1.2))
    sc.add_object(Object(Triangle([(4, -2, 3), (8, 0, 3.5), (4, 2, 3)]), Material((0.7, 0.7, 0.2), (0, 0, 0), 0.3)))
    sc.add_object(Object(Plane((0, 0, 0), (0, 0, 1)), Material.colored((0.5, 0.5, 0.5))))
    sc.add_object(Object(Triangle([(5, -3, 0.1), (5, 3, 0.1), (7, 1, 2)]), Material((0.2, 0.7, 0.9), (0, 0, 0), 1.0)))
    sc.add_object(Object(Sphere((6, 0, 9), 4), Material.light((3, 3, 3))))
    sc.add_object(Object(Triangle([(0, 0, 5), (0, 0, 5), (0, 4, 5)]), Material.mirror()))  # degenerate: never hit
    rgb, ids, ref = _compare_triangle_scene(rtx, oracle, sc, 320, 200, seed=9)
    assert {0, 1, 2}.issubset(set(np.unique(ids).tolist())) and sc.last_stats.triangle_panics == ref["panics"] == 0
    # with max_bounces = 0 there is no bounce, hence no libm step: everything is bit-exact again
    sc.config = sc.config.with_max_bounces(0)
    _compare(rtx, oracle, sc, 320, 200, seed=9, spp_tol=0.0)


def test_small_terrain_brute_force(rtx, oracle, gpu):
    """S3's generator at a small grid (2 x 24^2 triangles + light): brute force vs oracle, all gates."""
    from rtx_b200 import scenes
    sc = scenes.terrain(grid=25, rays_per_pixel=1, max_bounces=4)
    rgb, ids, ref = _compare_triangle_scene(rtx, oracle, sc, 160, 90, seed=1)
    assert (ids != rtx.MISS_ID).mean() > 0.3


def test_random_spheres_small(rtx, oracle, gpu):
    """S2's generator at 300 spheres (global-memory scene path when it exceeds the smem budget)."""
    from rtx_b200 import scenes
    sc = scenes.random_spheres(n=300, rays_per_pixel=2, max_bounces=8)
    _compare(rtx, oracle, sc, 200, 120, seed=1)
    sc = scenes.random_spheres(n=3000, rays_per_pixel=1, max_bounces=3)  # > 64 KB blob → global path
    _compare(rtx, oracle, sc, 96, 54, seed=1)


def test_render_to_image_matches_oracle_quantiser(rtx, oracle, gpu):  # scene.rs:173-178
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=0)  # bit-exact radiance → bit-exact bytes
    img = sc.render_to_image(200, 120, seed=2)
    ref = oracle.render_scene(sc, 200, 120, seed=2)
    assert np.array_equal(img, oracle.to_rgb8(ref["rgb"]))
    assert img.max() == 255 and img[0].sum() == 0  # lights saturate; top row is black sky


def test_sharded_render_is_byte_identical(rtx, gpu):
    """N-shard image == 1-shard image byte for byte (one GPU plays every rank; the N-rank NCCL path is
    covered by tests/test_dist_cpu.py and tests/test_dist_gpu.py)."""
    import torch
    from rtx_b200 import scenes
    from rtx_b200 import dist as rdist
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=8)
    w, h, seed = 200, 120, 5  # not multiples of the 16-pixel tile
    full, ids_full = sc.render(w, h, seed=seed, want_ids=True)
    for n in (2, 3, 8):
        rgb, ids, rays = rdist.render_all_shards_on_one_gpu(sc, w, h, seed, n)
        assert rgb.cpu().numpy().tobytes() == full.tobytes()
        assert np.array_equal(ids.cpu().numpy().view(np.uint32), ids_full)
        assert rays == sc.last_stats.rays


def test_golden_fixture(rtx, gpu):
    """Committed oracle output (tests/golden/make_golden.py): pins oracle and CUDA path against drift."""
    import os
    from rtx_b200 import scenes
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "sphere_demo_96x64_1spp_seed1.npz"))
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=8)
    rgb, ids = sc.render(96, 64, seed=1, want_ids=True)
    assert np.array_equal(ids, g["ids"]) and sc.last_stats.rays == int(g["rays"])
    assert np.abs(rgb - g["rgb"]).max() <= TOL_1SPP


def test_cpp_host_mirror(gpu):
    """The C++ mirror of the reference API (rust-raytracing_b200/host/rtx.hpp) drives the same library."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(__file__)), "rust-raytracing_b200", "lib", "rtx_host_test")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    print(out.stdout[-2000:], out.stderr[-2000:])
    assert out.returncode == 0 and "ALL OK" in out.stdout


def test_context_pool_kernel_equals_lane_kernel(rtx, gpu):
    """The two schedulers of the same arithmetic (rtx_render_pool.cu: 64 pixel contexts per warp in three state
    masks; rtx_render.cu: one pixel per lane) must produce the same bytes, ids, and counters — on spheres,
    planes and triangles, at ragged sizes, at 0 spp, at depth 0, and when sharded."""
    from rtx_b200 import Camera, Config, Material, Object, Plane, Scene, Sphere, Triangle, scenes
    BR, LANE = rtx.RTX_ACCEL_BRUTE, rtx.RTX_ACCEL_BRUTE | rtx.RTX_FLAG_LANE_KERNEL
    cases = []
    # (4, 8, 640, 360): more pixels than resident contexts, so both ends of the two-granularity queue are used
    for spp, depth, w, h in [(4, 8, 333, 77), (1, 0, 17, 5), (3, 2, 1, 1), (0, 8, 40, 30), (16, 8, 256, 144),
                             (4, 8, 640, 360), (5, 3, 700, 301)]:
        cases.append((scenes.sphere_demo(rays_per_pixel=spp, max_bounces=depth), w, h))
    tri = Scene(Config().with_rays_per_pixel(3).with_max_bounces(4), Camera((0, 0, 1), (1, 0, 0.2), 1.2))
    tri.add_object(Object(Triangle([(4, -2, 3), (8, 0, 3.5), (4, 2, 3)]), Material((0.7, 0.7, 0.2), (0, 0, 0), 0.3)))
    tri.add_object(Object(Plane((0, 0, 0), (0, 0, 1)), Material.colored((0.5, 0.5, 0.5))))
    tri.add_object(Object(Triangle([(5, -3, 0.1), (5, 3, 0.1), (7, 1, 2)]), Material((0.2, 0.7, 0.9), (0, 0, 0), 1.0)))
    tri.add_object(Object(Sphere((6, 0, 9), 4), Material.light((3, 3, 3))))
    cases.append((tri, 320, 200))
    cases.append((scenes.random_spheres(n=100, rays_per_pixel=2, max_bounces=8), 200, 120))  # 7 KB blob
    empty = Scene(Config().with_rays_per_pixel(2), Camera((0, 0, 0), (1, 0, 0), 1.0))
    for sc, w, h in cases:
        a, ia = sc.render(w, h, seed=5, flags=BR, want_ids=True)
        sa = (sc.last_stats.rays, sc.last_stats.paths, sc.last_stats.hits, sc.last_stats.prims_tested)
        b, ib = sc.render(w, h, seed=5, flags=LANE, want_ids=True)
        sb = (sc.last_stats.rays, sc.last_stats.paths, sc.last_stats.hits, sc.last_stats.prims_tested)
        assert a.tobytes() == b.tobytes() and np.array_equal(ia, ib) and sa == sb, (w, h, sa, sb)
        # both queue granularities of the pool kernel: whole pixels per context / one (pixel, sample) per context
        for q in (rtx.RTX_FLAG_PIXEL_QUEUE, rtx.RTX_FLAG_SAMPLE_QUEUE):
            c, ic = sc.render(w, h, seed=5, flags=BR | q, want_ids=True)
            sq = (sc.last_stats.rays, sc.last_stats.paths, sc.last_stats.hits, sc.last_stats.prims_tested)
            assert c.tobytes() == b.tobytes() and np.array_equal(ic, ib) and sq == sb, (w, h, q, sq, sb)
    a = empty.render(33, 9, seed=1)
    assert not a.any()


def test_pinned_host_buffers_are_written_by_the_kernel(rtx, gpu):
    """rtx_scene_render with pinned, device-mapped caller buffers stores pixels from the render kernel itself
    (no device frame, no D2H copy); pageable buffers go through the device frame.  Same bytes either way — pool
    kernel with both queues, lane kernel, LBVH, ids-only."""
    import torch
    from rtx_b200 import scenes
    cases = [(scenes.sphere_demo(rays_per_pixel=8, max_bounces=8), 333, 177, rtx.RTX_ACCEL_AUTO),
             (scenes.sphere_demo(rays_per_pixel=8, max_bounces=8), 333, 177, rtx.RTX_FLAG_SAMPLE_QUEUE),
             (scenes.sphere_demo(rays_per_pixel=2, max_bounces=3), 200, 90, rtx.RTX_FLAG_LANE_KERNEL),
             (scenes.random_spheres(n=500, rays_per_pixel=2, max_bounces=4), 240, 135, rtx.RTX_ACCEL_LBVH)]
    for sc, w, h, fl in cases:
        a, ia = sc.render(w, h, seed=9, flags=fl, want_ids=True)  # numpy: pageable
        pin = torch.full((h, w, 3), -1.0, dtype=torch.float64).pin_memory()
        pin_ids = torch.full((h, w), 12345, dtype=torch.int32).pin_memory()
        b, ib = sc.render(w, h, seed=9, flags=fl, out=pin.numpy(), out_ids=pin_ids.numpy().view(np.uint32))
        assert a.tobytes() == b.tobytes() and np.array_equal(ia, ib), (w, h, fl)
        ids_only = torch.full((h, w), 12345, dtype=torch.int32).pin_memory()
        st = rtx.Stats()
        hnd = sc._sync()
        import ctypes as C
        rtx._abi.check(rtx._abi.load().rtx_scene_render(hnd, w, h, 9, fl | rtx.RTX_FLAG_NO_RADIANCE, None,
                                                        ids_only.numpy().ctypes.data, C.byref(st)), hnd)
        assert np.array_equal(ids_only.numpy().view(np.uint32), ia)


def test_f32_sphere_cull_is_conservative(rtx, oracle, gpu):
    """The scan's f32 pre-test (rtx_trace.cuh, SphereCull) may only drop spheres the exact f64 test drops.  Scenes
    built to stress its error margin: grazing rays, huge and tiny coordinates, a camera far from the spheres,
    spheres at 1e19 (f32 overflow range of the squared terms) and 1e-25 (f32 underflow), a non-finite sphere."""
    from rtx_b200 import Camera, Config, Material, Object, Scene, Sphere
    def build(scale, cam_dist, extra=None):
        cfg = Config().with_rays_per_pixel(2).with_max_bounces(3)
        cfg = cfg.with_focal_length(10.0 * scale).with_focal_offset(1e-4 * scale).with_non_focal_offset(1e-4 * scale)
        sc = Scene(cfg,
                   Camera((-cam_dist * scale, 0.1 * scale, 0.2 * scale), (1, 0, 0), 0.02 if cam_dist > 100 else 1.0))
        for i, (x, y, z, r) in enumerate([(6, 0, 1, 1), (6, 2.0000001, 1, 1), (9, -1, 0.5, 1.5), (4, 1, -0.2, 0.25),
                                          (12, 0, 6, 3), (7, -3, 2, 0.999999)]):
            m = Material.light((2, 2, 2)) if i == 4 else Material((0.8, 0.5 + 0.05 * i, 0.3), (0, 0, 0), 0.3 * (i % 3))
            sc.add_object(Object(Sphere((x * scale, y * scale, z * scale), r * scale), m))
        if extra:
            sc.add_object(Object(Sphere(*extra), Material.colored((0.5, 0.5, 0.5))))
        return sc
    cases = [build(1.0, 0), build(1e6, 0), build(1e-6, 0), build(1.0, 5000.0), build(1e19, 0), build(1e-25, 0),
             build(1.0, 0, ((float("inf"), 0.0, 0.0), 1.0)), build(1.0, 0, ((3e38, 0.0, 0.0), 1.0)),
             build(1.0, 0, ((5.0, 0.0, float("nan")), 1.0))]
    for k, sc in enumerate(cases):
        rgb, ids = sc.render(160, 90, seed=3, flags=rtx.RTX_ACCEL_BRUTE, want_ids=True)
        ref = oracle.render_scene(sc, 160, 90, seed=3)
        assert np.array_equal(ids, ref["ids"]), k
        assert rgb.tobytes() == ref["rgb"].tobytes(), k
        assert sc.last_stats.rays == ref["rays"], k
        if k in (0, 1, 2, 3, 4, 5):
            assert (ids != 0xFFFFFFFF).any(), k  # the spheres are actually in view


def test_c2_full_size_frame_equals_oracle(rtx, oracle, gpu):
    """BASELINE.json configs[1] at full size — S1, 1920x1080, 64 spp, max depth 8: every byte of the frame, every
    primary hit id and the ray count equal the CPU oracle's (245 M rays; the oracle takes a few seconds on the
    box's host threads)."""
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=64, max_bounces=8)
    rgb, ids = sc.render(1920, 1080, seed=1, want_ids=True)
    ref = oracle.render_scene(sc, 1920, 1080, seed=1)
    assert np.array_equal(ids, ref["ids"])
    assert sc.last_stats.rays == ref["rays"] == 245087388
    assert float(np.abs(rgb - ref["rgb"]).max()) <= 1e-4 and psnr(rgb, ref["rgb"]) >= PSNR_MIN  # the north-star gates
    assert rgb.tobytes() == ref["rgb"].tobytes()  # the bar this repo holds: bit-exact
    # the same frame under the pixel queue: a pinned caller buffer (auto: pixels stored over PCIe by the kernel)
    # and forced on a device frame
    import torch
    pin = torch.full((1080, 1920, 3), -1.0, dtype=torch.float64).pin_memory()
    b = sc.render(1920, 1080, seed=1, out=pin.numpy())
    assert b.tobytes() == ref["rgb"].tobytes()
    c = sc.render(1920, 1080, seed=1, flags=rtx.RTX_FLAG_PIXEL_QUEUE)
    assert c.tobytes() == ref["rgb"].tobytes()


@pytest.mark.parametrize("w,h", [(2560, 1440), (3840, 2160), (3001, 1999)])
def test_banded_host_output_equals_oracle(rtx, oracle, gpu, w, h):
    """A large frame rendered into pinned host memory goes out in 2-4 bands of tile rows (sample queue + ordered
    sum into the device frame, copy engine to the caller's buffer while the next band renders).  Every byte,
    primary hit id and counter equals the oracle's, ragged borders included."""
    import torch
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=8)
    pin = torch.full((h, w, 3), -1.0, dtype=torch.float64).pin_memory()
    pin_ids = torch.full((h, w), 12345, dtype=torch.int32).pin_memory()
    rgb, ids = sc.render(w, h, seed=21, out=pin.numpy(), out_ids=pin_ids.numpy().view(np.uint32))
    st = (sc.last_stats.rays, sc.last_stats.paths, sc.last_stats.kernel_launches)
    ref = oracle.render_scene(sc, w, h, seed=21)
    assert st[2] >= 4, st  # at least two bands x (render + ordered sum)
    assert st[0] == ref["rays"] and st[1] == w * h * 4
    assert np.array_equal(ids, ref["ids"])
    assert rgb.tobytes() == ref["rgb"].tobytes()


def test_queue_plan_selection_and_launch_count(rtx, gpu):
    """rtx_scene_last_launches reports what the plan launched: the sample queue (render kernel + ordered-sum pass)
    whenever rays_per_pixel >= 4, the pixel queue (one kernel) below that, when forced, and for ids-only renders;
    LBVH scenes always one kernel."""
    import torch
    from rtx_b200 import scenes
    dev = torch.device("cuda", 0)
    w, h = 320, 200
    rgb = torch.empty((h, w, 3), dtype=torch.float64, device=dev)
    ids = torch.empty((h, w), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream

    def launches(sc, flags, want_rgb=True):
        sc.render_device(w, h, 1, flags, None, rgb.data_ptr() if want_rgb else None, ids.data_ptr(), None, st)
        torch.cuda.synchronize()
        return sc.last_device_launches

    s8 = scenes.sphere_demo(rays_per_pixel=8, max_bounces=4)
    s2 = scenes.sphere_demo(rays_per_pixel=2, max_bounces=4)
    assert launches(s8, rtx.RTX_ACCEL_AUTO) == 2
    assert launches(s8, rtx.RTX_FLAG_PIXEL_QUEUE) == 1
    assert launches(s8, rtx.RTX_FLAG_LANE_KERNEL) == 1
    assert launches(s8, rtx.RTX_FLAG_NO_RADIANCE, want_rgb=False) == 1
    assert launches(s2, rtx.RTX_ACCEL_AUTO) == 1
    assert launches(s2, rtx.RTX_FLAG_SAMPLE_QUEUE) == 2
    big = scenes.random_spheres(n=300, rays_per_pixel=8, max_bounces=2)
    assert launches(big, rtx.RTX_ACCEL_LBVH) == 1


def test_device_bands_compose_the_frame(rtx, gpu):
    """rtx_scene_render_device_band: the bands of a frame, rendered one by one into the same buffers, give the bytes of
    one whole-frame call — for the whole image and for every shard of a 3-way split (in place and compact), pool
    kernel under both queue plans and the LBVH lane kernel; rows outside a band are left alone."""
    import torch
    from rtx_b200 import scenes
    from rtx_b200 import dist as rdist
    dev = torch.device("cuda", 0)
    st = torch.cuda.current_stream(dev).cuda_stream
    w, h = 333, 211  # 14 tile rows, ragged in both directions
    cases = [(scenes.sphere_demo(rays_per_pixel=5, max_bounces=6), rtx.RTX_ACCEL_AUTO),
             (scenes.sphere_demo(rays_per_pixel=5, max_bounces=6), rtx.RTX_FLAG_PIXEL_QUEUE),
             (scenes.random_spheres(n=400, rays_per_pixel=2, max_bounces=3), rtx.RTX_ACCEL_LBVH)]
    for sc, fl in cases:
        for world in (1, 3):
            for rank in range(world):
                for in_place in ((True,) if world == 1 else (True, False)):
                    f = fl | (rtx.RTX_FLAG_SHARD_IN_PLACE if in_place else 0)
                    shard = (rank, world) if world > 1 else None
                    if in_place:
                        a = torch.full((h, w, 3), -7.0, dtype=torch.float64, device=dev)
                        ia = torch.full((h, w), -7, dtype=torch.int32, device=dev)
                    else:
                        a, ia = rdist.alloc_shard_buffers(w, h, world, dev, want_ids=True)
                        a.fill_(-7.0), ia.fill_(-7)
                    b, ib = a.clone(), ia.clone()
                    ca = torch.zeros(8, dtype=torch.int64, device=dev)
                    cb = torch.zeros(8, dtype=torch.int64, device=dev)
                    sc.render_device(w, h, 9, f, shard, a.data_ptr(), ia.data_ptr(), ca.data_ptr(), st)
                    for (r0, n) in ((0, 5), (5, 1), (6, 0), (6, 7), (13, 40)):
                        sc.render_device_band(w, h, 9, f, shard, r0, n, b.data_ptr(), ib.data_ptr(), cb.data_ptr(), st)
                    torch.cuda.synchronize()
                    assert torch.equal(a, b) and torch.equal(ia, ib), (fl, world, rank, in_place)
                    assert torch.equal(ca[:4], cb[:4]), (fl, world, rank, in_place)
    # a single band touches nothing outside its rows
    sc = cases[0][0]
    c = torch.full((h, w, 3), -7.0, dtype=torch.float64, device=dev)
    sc.render_device_band(w, h, 9, rtx.RTX_ACCEL_AUTO, None, 2, 3, c.data_ptr(), None, None, st)
    torch.cuda.synchronize()
    assert bool((c[:32] == -7.0).all()) and bool((c[80:] == -7.0).all()) and bool((c[32:80] != -7.0).all())


def test_two_renders_of_one_scene_on_two_streams(rtx, oracle, gpu):
    """Scene::render is `&self` (scene.rs:144): renders of one scene may be issued concurrently.  The scene owns device
    scratch (sample-queue slots, work counters), so the library orders such renders on the device with an event;
    both frames must equal the oracle's, whichever stream or thread issued them."""
    import threading
    import torch
    from rtx_b200 import scenes
    dev = torch.device("cuda", gpu)
    w, h = 640, 360
    sc = scenes.sphere_demo(rays_per_pixel=8, max_bounces=6)
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    out = [torch.full((h, w, 3), float("nan"), dtype=torch.float64, device=dev) for _ in range(4)]
    torch.cuda.synchronize(dev)
    # one thread, two streams, four frames back to back without a host synchronisation
    for i in range(4):
        sc.render_device(w, h, 20 + i, rtx.RTX_FLAG_SAMPLE_QUEUE, None, out[i].data_ptr(), None, None,
                         streams[i & 1].cuda_stream)
    torch.cuda.synchronize(dev)
    refs = [oracle.render_scene(sc, w, h, seed=20 + i)["rgb"] for i in range(4)]
    for i in range(4):
        assert float(np.abs(out[i].cpu().numpy() - refs[i]).max()) <= 1e-4, f"frame {i} (stream {i & 1})"
    first = [o.cpu().numpy().copy() for o in out]
    # two threads, a stream each
    for o in out:
        o.fill_(float("nan"))
    torch.cuda.synchronize(dev)

    def work(t):
        for i in (t, t + 2):
            sc.render_device(w, h, 20 + i, rtx.RTX_FLAG_SAMPLE_QUEUE, None, out[i].data_ptr(), None, None,
                             streams[t].cuda_stream)
    th = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    torch.cuda.synchronize(dev)
    for i in range(4):
        assert out[i].cpu().numpy().tobytes() == first[i].tobytes(), f"threaded frame {i}"


def test_reference_panic_is_reported(rtx, oracle, gpu):
    """A triangle whose elimination trips an assert_eq! of the reference (triangle.rs:76-94; here: an infinite
    coordinate makes lgs1.x / lgs1.x NaN) panics the reference's row thread and aborts Scene::render.  The CUDA path
    rejects the hit, counts it, completes the image — and rtx_scene_render says so with RTX_ERR_REFERENCE_PANIC."""
    import ctypes as C
    from rtx_b200 import Material, Object, Triangle, scenes
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=2)
    w, h = 64, 48
    clean = sc.render(w, h, seed=3)
    sc.add_object(Object(Triangle([(float("inf"), 0, 0), (5, 1, 0), (5, 0, 1)]), Material.colored((1, 1, 1))))
    with pytest.raises(rtx.RtxError) as e:
        sc.render(w, h, seed=3)
    assert e.value.status == rtx._abi.RTX_ERR_REFERENCE_PANIC and "assert_eq" in str(e.value)
    # the outputs and statistics are complete all the same: through the C ABI directly
    lib, hnd = rtx._abi.load(), sc._sync()
    rgb = np.full((h, w, 3), np.nan)
    st = rtx._abi.Stats()
    status = lib.rtx_scene_render(hnd, w, h, 3, 0, rgb.ctypes.data, None, C.byref(st))
    assert status == rtx._abi.RTX_ERR_REFERENCE_PANIC and st.triangle_panics > 0 and st.pixels == w * h
    ref = oracle.render_scene(sc, w, h, seed=3)
    assert ref["panics"] == st.triangle_panics
    assert rgb.tobytes() == clean.tobytes()  # the rejected triangle changes nothing


def test_rng_row_mode_is_the_oracles_row_mode(rtx, oracle, gpu):
    """RTX_FLAG_RNG_ROW: one CUDA thread per image row consuming one stream exactly as a reference row thread does
    (scene.rs:149-161; every draw, also the bounce directions nothing traces).  Against the oracle's RNG_ROW mode:
    hit ids and ray counts equal, radiance within the 1-spp gate (and bit-equal where no path bounced)."""
    from rtx_b200 import scenes
    for sc, flags, (w, h) in ((scenes.sphere_demo(rays_per_pixel=2, max_bounces=4), rtx.RTX_ACCEL_AUTO, (96, 54)),
                              (scenes.sphere_demo(rays_per_pixel=1, max_bounces=0), rtx.RTX_ACCEL_AUTO, (33, 17)),
                              (scenes.random_spheres(n=200, rays_per_pixel=1, max_bounces=3), rtx.RTX_ACCEL_LBVH, (48, 32))):
        rgb, ids = sc.render(w, h, seed=7, flags=flags | rtx.RTX_FLAG_RNG_ROW, want_ids=True)
        ref = oracle.render_scene(sc, w, h, seed=7, rng_mode=oracle.RNG_ROW)
        assert np.array_equal(ids, ref["ids"])
        assert sc.last_stats.rays == ref["rays"]
        assert sc.last_stats.kernel_launches == 1 or flags == rtx.RTX_ACCEL_LBVH  # (+ the LBVH build's launches)
        assert float(np.abs(rgb - ref["rgb"]).max()) <= TOL_1SPP
        assert rgb.tobytes() == ref["rgb"].tobytes()  # the device sin/cos are glibc's bit for bit
        # and it is NOT the default image (other seeding), while primary hit ids may differ only through jitter
        assert sc.render(w, h, seed=7, flags=flags).tobytes() != rgb.tobytes() or sc.config.max_bounces == 0
    with pytest.raises(rtx.RtxError):
        import torch
        t = torch.empty((17, 33, 3), dtype=torch.float64, device="cuda")
        sc.render_device(33, 17, 1, rtx.RTX_FLAG_RNG_ROW, (0, 2), t.data_ptr(), None, None, 0)


@pytest.mark.parametrize("n_sph", [256, 1024, 1025, 2500])
def test_chunked_scan_of_a_large_sphere_scene_equals_the_oracle(rtx, oracle, gpu, n_sph):
    """A scanned scene beyond the 64 KB shared-memory stage: the lane kernel streams the sphere arrays through shared
    memory in chunks of 1024 (two TMA-fed buffers per CTA, rtx_render.cu closest_chunked).  Chunk counts 1, 1, 2 (one
    sphere in the second) and 3 (ragged last chunk), with planes and triangles (read in place) mixed in, against the
    oracle: ids and ray counts equal, radiance within the 1-spp gate; and against the LBVH render, byte for byte."""
    from rtx_b200 import Material, Object, Plane, Triangle, scenes
    sc = scenes.random_spheres(n=n_sph, rays_per_pixel=2, max_bounces=4)
    sc.add_object(Object(Plane((0, 0, -45), (0, 0, 1)), Material((0.7, 0.7, 0.7), (0, 0, 0), 0.9)))
    sc.add_object(Object(Triangle([(20, -10, -10), (20, 10, -10), (20, 0, 12)]), Material((0.9, 0.4, 0.2), (0.1, 0.1, 0.1), 0.4)))
    w, h = 160, 90
    rgb, ids = sc.render(w, h, seed=5, flags=rtx.RTX_ACCEL_BRUTE, want_ids=True)
    assert sc.last_stats.accel_used == rtx.RTX_ACCEL_BRUTE and sc.last_stats.prims_tested == sc.last_stats.rays * (n_sph + 2)
    ref = oracle.render_scene(sc, w, h, seed=5)
    assert np.array_equal(ids, ref["ids"]) and sc.last_stats.rays == ref["rays"]
    assert float(np.abs(rgb - ref["rgb"]).max()) <= TOL_1SPP
    via_tree = sc.render(w, h, seed=5, flags=rtx.RTX_ACCEL_LBVH)
    assert via_tree.tobytes() == rgb.tobytes()
