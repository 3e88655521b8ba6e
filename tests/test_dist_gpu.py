"""GPU, world_size 2, NCCL: the real multi-GPU path — each rank renders its interleaved tiles, one gather to
rank 0, de-interleave kernel — must equal the single-GPU image byte for byte.  Needs 2 GPUs (skipped on 1)."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, w, h, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from rtx_b200 import dist as rdist
        from rtx_b200 import scenes
        sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=8, device=rank)
        rgb, ids, counters = rdist.render_distributed(sc, w, h, seed=3)
        total = counters.clone()
        dist.all_reduce(total)
        # the same image again, gather fused into the render kernel: peers store into rank 0's frame (NVLink)
        rgb2, ids2, counters2, frame = rdist.render_distributed_p2p(sc, w, h, seed=3)
        total2 = counters2.clone()
        dist.all_reduce(total2)
        if rank == 0:  # views of the frame: keep copies, the frame is poisoned and reused below
            rgb2, ids2 = rgb2.clone(), ids2.clone()
        # and handed over to pinned host memory band by band (PeerFrame.render_to_host): other seeds, so that a stale
        # or missed band cannot compare equal, into a poisoned frame and a poisoned host buffer; two frames back to
        # back without a host synchronisation in between (the second must not tear the first one's copy-out)
        pin = [torch.full((h, w, 3), -1.0, dtype=torch.float64).pin_memory() if rank == 0 else None for _ in range(2)]
        frame.poison()
        frame.render_to_host(sc, 11, pin[0], bands=3)
        frame.render_to_host(sc, 12, pin[1], bands=2)
        torch.cuda.synchronize()
        # a device-resident frame again after the host frames: double buffering keeps it apart from the copies
        frame.poison()
        frame.render(sc, 13)
        torch.cuda.synchronize()
        if rank == 0:
            rgb4 = frame.tensors()[0].cpu().numpy()
            full, ids_full = sc.render(w, h, seed=3, want_ids=True)
            ok = rgb.cpu().numpy().tobytes() == full.tobytes()
            ok = ok and np.array_equal(ids.cpu().numpy().view(np.uint32), ids_full)
            ok = ok and int(total[0].item()) == sc.last_stats.rays
            ok2 = rgb2.cpu().numpy().tobytes() == full.tobytes()
            ok2 = ok2 and np.array_equal(ids2.cpu().numpy().view(np.uint32), ids_full)
            ok2 = ok2 and int(total2[0].item()) == sc.last_stats.rays
            ok3 = pin[0].numpy().tobytes() == sc.render(w, h, seed=11).tobytes()
            ok3 = ok3 and pin[1].numpy().tobytes() == sc.render(w, h, seed=12).tobytes()
            ok3 = ok3 and rgb4.tobytes() == sc.render(w, h, seed=13).tobytes()
            q.put((bool(ok), bool(ok2), bool(ok3)))
        dist.barrier()
        frame.close()
        # no rank-0 funnel: one pinned host frame in shared memory, every rank delivers its own rows over its own PCIe
        # link (rtx_scene_render_rows_to_host); two frames back to back, then the completion barrier
        from rtx_b200 import HostFrame
        name = f"/rtx-test-dist-{port}"
        hf = HostFrame(name, w, h, create=True) if rank == 0 else None
        dist.barrier()
        if rank != 0:
            hf = HostFrame(name, w, h, create=False)
        if rank == 0:
            hf.array[:] = -1.0
        dist.barrier()
        st = torch.cuda.current_stream().cuda_stream
        ok4 = True
        for seed in (21, 22):
            sc.render_rows_to_host(w, h, seed, 0, (rank, world), hf.ptr, None, None, st)
            torch.cuda.synchronize()
            dist.barrier()
            if rank == 0:
                ok4 = ok4 and hf.array.tobytes() == sc.render(w, h, seed=seed).tobytes()
            dist.barrier()
        hf.close()
        if rank == 0:
            q.put(bool(ok4))
    finally:
        dist.destroy_process_group()


def test_two_gpu_render_equals_one_gpu(rtx, gpu):
    if rtx.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1000, 562, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert q.get(timeout=5) == (True, True, True)  # (NCCL gather path, fused peer-store path, banded hand-over to the host)
    assert q.get(timeout=5) is True  # shared host frame, every rank delivering its own rows


def test_multi_handle_two_gpus_equals_one_gpu(rtx, oracle, gpu):
    """rtx_multi_*: both GPUs of this process behind one handle, no process group; the image is the one-GPU image and
    the oracle's."""
    if rtx.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    import torch
    from rtx_b200 import MultiScene, scenes
    w, h = 1000, 562
    sc = scenes.sphere_demo(rays_per_pixel=4, max_bounces=8)
    m = MultiScene(sc, [0, 1])
    pin = torch.full((h, w, 3), -1.0, dtype=torch.float64).pin_memory()
    for seed in (5, 6, 7):
        got, ids = m.render(w, h, seed=seed, out=pin.numpy(), want_ids=True)
        ref, ref_ids = sc.render(w, h, seed=seed, want_ids=True)
        assert got.tobytes() == ref.tobytes() and np.array_equal(ids, ref_ids), seed
        assert m.last_stats.rays == sc.last_stats.rays
    assert float(np.abs(got - oracle.render_scene(sc, w, h, seed=7)["rgb"]).max()) <= 1e-4
    pageable = m.render(w, h, seed=7)  # a pageable caller buffer works too (slower copies)
    assert pageable.tobytes() == ref.tobytes()
    m.close()
