"""Row sharding (RTX_FLAG_SHARD_ROWS), delivery of a shard's rows to a host frame (rtx_scene_render_rows_to_host),
the shared host frame (rtx_host_frame_*) and the in-process multi-GPU handle (rtx_multi_*), on ONE GPU playing every
rank in turn; the 2-GPU versions are in test_dist_gpu.py.  Gate: the bytes of the one-GPU whole-frame render, which
other tests compare with the oracle — here too, once per scene."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _scenes(rtx):
    from rtx_b200 import scenes
    return [("pool kernel, sample queue", scenes.sphere_demo(rays_per_pixel=6, max_bounces=5), rtx.RTX_ACCEL_AUTO),
            ("pool kernel, pixel queue", scenes.sphere_demo(rays_per_pixel=2, max_bounces=5), rtx.RTX_ACCEL_AUTO),
            ("lane kernel, primary only", scenes.sphere_demo(rays_per_pixel=1, max_bounces=0), rtx.RTX_ACCEL_AUTO),
            ("LBVH", scenes.random_spheres(n=300, rays_per_pixel=2, max_bounces=3), rtx.RTX_ACCEL_LBVH)]


@pytest.mark.parametrize("w,h", [(200, 120), (67, 45), (33, 3)])
def test_row_shards_compose_the_frame(rtx, oracle, gpu, w, h):
    """Strip s (4 pixel rows) belongs to shard s % count: every shard of 1, 2, 3 and 5, rendered in place into one
    frame — whole, and band by band — gives the bytes of the whole-frame render, for every kernel."""
    import torch
    dev = torch.device("cuda", gpu)
    st = torch.cuda.current_stream(dev).cuda_stream
    for name, sc, accel in _scenes(rtx):
        want, want_ids = sc.render(w, h, seed=4, flags=accel, want_ids=True)
        if name.startswith("pool kernel, sample"):
            assert float(np.abs(want - oracle.render_scene(sc, w, h, seed=4)["rgb"]).max()) <= 1e-4
        for count in (1, 2, 3, 5):
            a = torch.full((h, w, 3), float("nan"), dtype=torch.float64, device=dev)
            ia = torch.full((h, w), -7, dtype=torch.int32, device=dev)
            b, ib = a.clone(), ia.clone()
            cnt = torch.zeros(8, dtype=torch.int64, device=dev)
            f = accel | rtx.RTX_FLAG_SHARD_ROWS
            for r in range(count):
                sc.render_device(w, h, 4, f, (r, count), a.data_ptr(), ia.data_ptr(), cnt.data_ptr(), st)
                for (r0, n) in ((0, 1), (1, 2), (3, 100)):
                    sc.render_device_band(w, h, 4, f, (r, count), r0, n, b.data_ptr(), ib.data_ptr(), None, st)
            torch.cuda.synchronize(dev)
            assert a.cpu().numpy().tobytes() == want.tobytes(), (name, count)
            assert np.array_equal(ia.cpu().numpy().view(np.uint32), want_ids), (name, count)
            assert b.cpu().numpy().tobytes() == want.tobytes() and np.array_equal(ib.cpu().numpy().view(np.uint32), want_ids)
            assert int(cnt[0].item()) == sc.last_stats.rays


@pytest.mark.parametrize("w,h", [(640, 360), (203, 131), (1920, 1080)])
def test_rows_to_host_every_rank_delivers_its_own_rows(rtx, gpu, w, h):
    import torch
    from rtx_b200 import scenes
    dev = torch.device("cuda", gpu)
    st = torch.cuda.current_stream(dev).cuda_stream
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=6)
    want, want_ids = sc.render(w, h, seed=9, want_ids=True)
    for count in (1, 3, 8):
        pin = torch.full((h, w, 3), -1.0, dtype=torch.float64).pin_memory()
        pid = torch.full((h, w), -7, dtype=torch.int32).pin_memory()
        cnt = torch.zeros(8, dtype=torch.int64, device=dev)
        for r in range(count):
            sc.render_rows_to_host(w, h, 9, 0, (r, count), pin.data_ptr(), pid.data_ptr(), cnt.data_ptr(), st)
        torch.cuda.synchronize(dev)
        assert pin.numpy().tobytes() == want.tobytes(), count
        assert np.array_equal(pin.numpy(), want) and np.array_equal(pid.numpy().view(np.uint32), want_ids)
        assert int(cnt[0].item()) == sc.last_stats.rays


def test_host_frame_and_multi_handle_on_one_gpu(rtx, gpu):
    """rtx_host_frame_create / open map the same pinned memory (here twice in one process); rtx_multi_* with a single
    device renders the one-GPU image through the worker-thread path."""
    import torch
    from rtx_b200 import HostFrame, MultiScene, scenes
    torch.cuda.set_device(gpu)
    w, h = 320, 200
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=6)
    want, want_ids = sc.render(w, h, seed=11, want_ids=True)
    name = f"/rtx-test-frame-{os.getpid()}"
    fa = HostFrame(name, w, h, create=True)
    try:
        with pytest.raises(rtx.RtxError):
            HostFrame(name, w, h, create=True)  # exclusive
        fb = HostFrame(name, w, h, create=False)
        fa.array[:] = -1.0
        st = torch.cuda.current_stream().cuda_stream
        sc.render_rows_to_host(w, h, 11, 0, (0, 2), fa.ptr, None, None, st)   # rank 0 through the creator's mapping
        sc.render_rows_to_host(w, h, 11, 0, (1, 2), fb.ptr, None, None, st)   # rank 1 through the opener's
        torch.cuda.synchronize()
        assert fa.array.tobytes() == want.tobytes() and fb.array.tobytes() == want.tobytes()
        fb.close()
    finally:
        fa.close()
    with pytest.raises(rtx.RtxError):
        HostFrame(name, w, h, create=False)  # unlinked by the creator's close
    m = MultiScene(sc, [gpu])
    for seed in (11, 12):
        got, ids = m.render(w, h, seed=seed, want_ids=True)
        ref, ref_ids = sc.render(w, h, seed=seed, want_ids=True)
        assert got.tobytes() == ref.tobytes() and np.array_equal(ids, ref_ids)
        assert m.last_stats.rays == sc.last_stats.rays and m.last_stats.pixels == w * h
    m.close()


def test_error_paths_of_the_multi_gpu_entry_points(rtx, gpu):
    """Bad arguments are refused with RTX_ERR_INVALID_ARGUMENT and a message, not with a crash or a silent no-op."""
    import ctypes as C
    import torch
    from rtx_b200 import HostFrame, MultiScene, scenes
    lib = rtx._abi.load()
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=2)
    h = sc._sync()
    pin = torch.empty((64, 64, 3), dtype=torch.float64).pin_memory()
    st = torch.cuda.current_stream().cuda_stream
    sh = rtx._abi.Shard(2, 2)  # index out of range
    assert lib.rtx_scene_render_rows_to_host(h, 64, 64, 1, 0, C.byref(sh), pin.data_ptr(), None, None, st) == rtx._abi.RTX_ERR_INVALID_ARGUMENT
    assert lib.rtx_scene_render_rows_to_host(h, 64, 64, 1, 0, None, None, None, None, st) == rtx._abi.RTX_ERR_INVALID_ARGUMENT
    assert lib.rtx_scene_render_rows_to_host(h, 0, 64, 1, 0, None, pin.data_ptr(), None, None, st) == rtx._abi.RTX_ERR_INVALID_ARGUMENT
    assert b"host_rgb" in lib.rtx_last_error(h) or b"size" in lib.rtx_last_error(h)
    # more shards than strips: the extra shards own nothing and the call is a no-op for them
    pin.fill_(-1.0)
    for r in range(40):
        sc.render_rows_to_host(64, 64, 3, 0, (r, 40), pin.data_ptr(), None, None, st)
    torch.cuda.synchronize()
    assert pin.numpy().tobytes() == sc.render(64, 64, seed=3).tobytes()
    with pytest.raises(rtx.RtxError):
        MultiScene(sc, [gpu, gpu])       # a device listed twice
    with pytest.raises(rtx.RtxError):
        MultiScene(sc, [rtx.device_count() + 3])  # no such device
    with pytest.raises(rtx.RtxError):
        MultiScene(sc, [])
    with pytest.raises(rtx.RtxError):
        HostFrame("/rtx-test-does-not-exist", 8, 8, create=False)
    assert lib.rtx_multi_render(None, 8, 8, 0, 0, pin.data_ptr(), None, None) == rtx._abi.RTX_ERR_INVALID_ARGUMENT
    assert lib.rtx_multi_device_count(None) == 0 and lib.rtx_multi_scene(None, 0) is None


def test_multi_render_from_two_threads(rtx, gpu):
    """rtx_multi_render is one frame at a time inside the library: two threads rendering different seeds through one
    handle each get their own frame and their own statistics."""
    import threading
    from rtx_b200 import MultiScene, scenes
    sc = scenes.sphere_demo(rays_per_pixel=3, max_bounces=5)
    w, h = 160, 100
    want = {seed: (sc.render(w, h, seed=seed), sc.last_stats.rays) for seed in (31, 32)}
    m = MultiScene(sc, [gpu])
    out = {}
    lock = threading.Lock()

    def work(seed):
        import ctypes as C
        for _ in range(6):
            rgb = np.empty((h, w, 3), dtype=np.float64)
            st = rtx._abi.Stats()
            rc = rtx._abi.load().rtx_multi_render(m._handle, w, h, seed, 0, rgb.ctypes.data, None, C.byref(st))
            with lock:
                out.setdefault(seed, []).append((rc, rgb.tobytes(), st.rays))
    th = [threading.Thread(target=work, args=(s,)) for s in (31, 32)]
    [t.start() for t in th]
    [t.join() for t in th]
    for seed in (31, 32):
        assert len(out[seed]) == 6
        for rc, raw, rays in out[seed]:
            assert rc == 0 and raw == want[seed][0].tobytes() and rays == want[seed][1]
    m.close()
