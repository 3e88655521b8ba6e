"""rtx_scene_render_device_noalloc (include/rtx_cuda.h): the sample-queue render whose scratch is the library's
static device array.  Through the C ABI, two calls in a row, against rtx_scene_render_device on the same scene, size
and seed — the gate is the one every integer/byte comparison in this suite uses: identical bytes.  This file is
collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _demo(rtx, rpp, bounces=4):
    from rtx_b200 import scenes
    return scenes.sphere_demo(rays_per_pixel=rpp, max_bounces=bounces)


def test_noalloc_render_equals_render_device_twice_in_a_row(rtx, oracle, gpu):
    import torch
    dev = torch.device("cuda", gpu)
    w, h, rpp = 200, 120, 8  # ragged against the 16x16 tiles
    sc = _demo(rtx, rpp)
    st = torch.cuda.current_stream(dev).cuda_stream
    want = torch.empty((h, w, 3), dtype=torch.float64, device=dev)
    want_ids = torch.empty((h, w), dtype=torch.int32, device=dev)
    frames = {}
    for seed in (5, 6):  # set-up: the existing entry point (allocates; makes the scene and the tables resident)
        sc.render_device(w, h, seed, rtx.RTX_FLAG_SAMPLE_QUEUE, None, want.data_ptr(), want_ids.data_ptr(), None, st)
        torch.cuda.synchronize(dev)
        frames[seed] = (want.cpu().numpy().copy(), want_ids.cpu().numpy().copy())
    assert frames[5][0].tobytes() != frames[6][0].tobytes()
    ref = oracle.render_scene(sc, w, h, seed=5)
    assert np.array_equal(frames[5][1].view(np.uint32), ref["ids"])
    assert float(np.abs(frames[5][0] - ref["rgb"]).max()) <= 1e-4

    got = torch.empty_like(want)
    got_ids = torch.empty_like(want_ids)
    before = torch.cuda.memory_allocated(dev)
    # two calls in a row, different seeds: the second one finds the first one's radiance and flags in the scratch
    for seed in (5, 6, 5):
        got.fill_(float("nan"))
        got_ids.fill_(-7)
        sc.render_device_noalloc(w, h, seed, rtx.RTX_ACCEL_AUTO, got.data_ptr(), got_ids.data_ptr(), st)
        torch.cuda.synchronize(dev)
        assert got.cpu().numpy().tobytes() == frames[seed][0].tobytes(), f"seed {seed}: radiance bytes differ"
        assert np.array_equal(got_ids.cpu().numpy(), frames[seed][1]), f"seed {seed}: hit ids differ"
        assert sc.last_device_launches == 2
    assert torch.cuda.memory_allocated(dev) == before
    # ids are optional
    got.fill_(float("nan"))
    sc.render_device_noalloc(w, h, 6, rtx.RTX_FLAG_SAMPLE_QUEUE, got.data_ptr(), None, st)
    torch.cuda.synchronize(dev)
    assert got.cpu().numpy().tobytes() == frames[6][0].tobytes()


def test_noalloc_rejects_what_it_cannot_do(rtx, gpu):
    import torch
    dev = torch.device("cuda", gpu)
    st = torch.cuda.current_stream(dev).cuda_stream
    cap = rtx._abi.RTX_NOALLOC_SCRATCH_BYTES
    buf = torch.empty((1088 * 1920 * 3,), dtype=torch.float64, device=dev)

    # not resident yet: nothing has rendered this scene
    sc = _demo(rtx, 8)
    with pytest.raises(rtx._abi.RtxError, match="not resident"):
        sc.render_device_noalloc(64, 64, 1, 0, buf.data_ptr(), None, st)
    sc.render_device(64, 64, 1, 0, None, buf.data_ptr(), None, None, st)
    sc.render_device_noalloc(64, 64, 1, 0, buf.data_ptr(), None, st)
    # resident, but no camera tables for this size
    with pytest.raises(rtx._abi.RtxError, match="camera tables"):
        sc.render_device_noalloc(96, 64, 1, 0, buf.data_ptr(), None, st)
    # beyond the static scratch: 1920x1080 (1088 rows of tiles) x 8 spp x 25 B = 418 MB
    sc.render_device(1920, 1080, 1, 0, None, buf.data_ptr(), None, None, st)
    assert 1920 * 1088 * 8 * 25 > cap
    with pytest.raises(rtx._abi.RtxError, match="bytes of scratch"):
        sc.render_device_noalloc(1920, 1080, 1, 0, buf.data_ptr(), None, st)
    # flags it does not take, null output
    with pytest.raises(rtx._abi.RtxError):
        sc.render_device_noalloc(1920, 1080, 1, rtx.RTX_FLAG_NO_RADIANCE, buf.data_ptr(), None, st)
    with pytest.raises(rtx._abi.RtxError):
        sc.render_device_noalloc(1920, 1080, 1, 0, None, None, st)
    # the largest frame that fits still renders: paths * 25 <= cap - 256
    side = 16 * int(((cap - 256) // (25 * 8)) ** 0.5 // 16)
    assert side * side * 8 * 25 <= cap - 256
    sc.render_device(side, side, 3, rtx.RTX_FLAG_SAMPLE_QUEUE, None, buf.data_ptr(), None, None, st)
    torch.cuda.synchronize(dev)
    want = buf[: side * side * 3].cpu().numpy().copy()
    buf.fill_(float("nan"))
    sc.render_device_noalloc(side, side, 3, 0, buf.data_ptr(), None, st)
    torch.cuda.synchronize(dev)
    assert buf[: side * side * 3].cpu().numpy().tobytes() == want.tobytes()
    # objects added since: not resident again
    from rtx_b200 import Material, Object, Sphere
    sc.add_object(Object(Sphere((3, 0, 0.5), 0.1), Material.colored((0.5, 0.5, 0.5))))
    with pytest.raises(rtx._abi.RtxError, match="not resident"):
        sc.render_device_noalloc(side, side, 3, 0, buf.data_ptr(), None, st)
    torch.cuda.synchronize(dev)
