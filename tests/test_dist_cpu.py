"""CPU, world_size 2, gloo: the host logic of the multi-GPU path — tile ownership, compact shard layout,
gather to rank 0 — without a GPU.  Pixels come from the oracle (cut into each rank's compact tiles by a
test-local helper); the product's de-interleave is a CUDA kernel, so its index math is checked here
against a numpy restatement, and byte-for-byte on the GPU in test_gpu_render.py / test_dist_gpu.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
T = 16


def compact_tiles(img, rank, world, slots):
    """Full image [h][w][c] → rank's compact buffer [slots][16][16][c] (zero padded)."""
    h, w = img.shape[:2]
    tx_n, ty_n = (w + T - 1) // T, (h + T - 1) // T
    out = np.zeros((slots, T, T) + img.shape[2:], dtype=img.dtype)
    for gt in range(rank, tx_n * ty_n, world):
        ty, tx = divmod(gt, tx_n)
        blk = img[ty * T:(ty + 1) * T, tx * T:(tx + 1) * T]
        out[gt // world, :blk.shape[0], :blk.shape[1]] = blk
    return out


def deinterleave_numpy(g, w, h, world):
    """Restates rtx_deinterleave_tiles' index math: gathered [world][slots][16][16][c] → [h][w][c]."""
    tx_n = (w + T - 1) // T
    out = np.empty((h, w) + g.shape[4:], dtype=g.dtype)
    for y in range(h):
        for x in range(w):
            gt = (y // T) * tx_n + (x // T)
            out[y, x] = g[gt % world, gt // world, y % T, x % T]
    return out


def _worker(rank, world, port, w, h, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        from rtx_b200 import dist as rdist
        from rtx_b200 import scenes, shard_tile_count
        sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=2)
        full = oracle.render_scene(sc, w, h, seed=4, nthreads=2)
        slots = rdist.shard_slots(w, h, world)
        mine = int(shard_tile_count(w, h, rank, world))
        assert mine <= slots
        rgb_c = torch.from_numpy(compact_tiles(full["rgb"], rank, world, slots))
        ids_c = torch.from_numpy(compact_tiles(full["ids"].view(np.int32), rank, world, slots))
        g_rgb = rdist.gather_shards(rgb_c)
        g_ids = rdist.gather_shards(ids_c)
        if rank == 0:
            assert tuple(g_rgb.shape) == (world, slots, T, T, 3)
            img = deinterleave_numpy(g_rgb.numpy(), w, h, world)
            ids = deinterleave_numpy(g_ids.numpy(), w, h, world)
            ok = img.tobytes() == full["rgb"].tobytes() and np.array_equal(ids.view(np.uint32), full["ids"])
            try:
                rdist.deinterleave(w, h, world, g_rgb, g_ids)
                ok = False  # must refuse CPU tensors: no CPU fallback
            except RuntimeError:
                pass
            q.put(ok)
        else:
            assert g_rgb is None and g_ids is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gather_reassembles_the_image():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 72, 40, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_tile_ownership_is_a_partition():
    sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
    from rtx_b200 import shard_tile_count
    for (w, h) in ((72, 40), (1920, 1080), (3840, 2160), (17, 16)):
        tiles = ((w + T - 1) // T) * ((h + T - 1) // T)
        for world in (1, 2, 4, 8):
            owned = [len(range(r, tiles, world)) for r in range(world)]
            assert owned == [int(shard_tile_count(w, h, r, world)) for r in range(world)]
            assert sum(owned) == tiles


def test_band_tiles_match_enumeration():
    """rtx_shard_band_tiles (the arithmetic behind rtx_scene_render_device_band): the local tiles of shard r of N whose
    image tile lies in the tile rows [first, first + count) — compared with plain enumeration, including empty bands,
    bands running past the image and degenerate shards; the bands of PeerFrame.band_rows tile every shard exactly."""
    import ctypes as C
    import types
    sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
    from rtx_b200 import _abi
    from rtx_b200.dist import PeerFrame
    lib = _abi.load()

    def band(w, h, r, n, r0, cnt):
        a, b = C.c_uint32(123), C.c_uint32(456)
        lib.rtx_shard_band_tiles(w, h, r, n, r0, cnt, C.byref(a), C.byref(b))
        return a.value, b.value

    for (w, h) in ((72, 40), (333, 211), (1920, 1080), (17, 16), (16, 1)):
        tx, ty = (w + T - 1) // T, (h + T - 1) // T
        for world in (1, 2, 3, 8):
            for r in range(world):
                mine = list(range(r, tx * ty, world))  # image tile index of local tile k
                for (r0, cnt) in ((0, ty), (0, 1), (1, 2), (ty - 1, 1), (ty - 1, 5), (ty, 3), (2, 0), (0, 10**6)):
                    lo, hi = min(r0, ty) * tx, min(r0 + cnt, ty) * tx
                    want = [k for k, t in enumerate(mine) if lo <= t < hi]
                    first, n = band(w, h, r, world, r0, cnt)
                    assert n == len(want), (w, h, world, r, r0, cnt)
                    assert n == 0 or first == want[0]
                # the bands a PeerFrame hands out cover the shard's tiles once, in order
                for bands in (1, 2, 4, 7, 100):
                    rows = PeerFrame.band_rows(types.SimpleNamespace(height=h), bands)
                    assert rows[0][0] == 0 and rows[-1][1] == ty
                    assert all(a[1] == b[0] for a, b in zip(rows, rows[1:]))
                    got, nxt = 0, 0
                    for (a, b) in rows:
                        first, n = band(w, h, r, world, a, b - a)
                        assert n == 0 or first == nxt
                        nxt += n
                        got += n
                    assert got == len(mine)
    assert band(64, 64, 3, 2, 0, 4) == (0, 0) and band(64, 64, 0, 0, 0, 4) == (0, 0)  # invalid shards: nothing


def test_row_strips_match_enumeration_and_compose_the_oracle_image():
    """RTX_FLAG_SHARD_ROWS arithmetic (rtx_shard_strip_count / rtx_shard_band_strips, host only): strip s of 4 pixel rows
    belongs to shard s % count as its local strip s // count; a band of tile rows is a contiguous range of local strips.
    Against plain enumeration; and the rows the shards own, cut from an oracle image and put back strip by strip exactly
    as rtx_scene_render_rows_to_host's 2-D copies place them (row = strip, pitch = count strips, ragged last strip
    apart), give the image."""
    import ctypes as C
    sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
    sys.path.insert(0, ROOT)
    from oracle import oracle
    from rtx_b200 import _abi, scenes
    lib = _abi.load()

    def band(h, r, n, r0, cnt):
        a, b = C.c_uint32(123), C.c_uint32(456)
        lib.rtx_shard_band_strips(h, r, n, r0, cnt, C.byref(a), C.byref(b))
        return a.value, b.value

    for h in (1, 3, 4, 5, 40, 131, 1080):
        strips, ty = (h + 3) // 4, (h + T - 1) // T
        for world in (1, 2, 3, 8):
            owned = []
            for r in range(world):
                mine = list(range(r, strips, world))
                assert lib.rtx_shard_strip_count(h, r, world) == len(mine)
                owned += mine
                for (r0, cnt) in ((0, ty), (0, 1), (1, 2), (ty - 1, 1), (ty - 1, 5), (ty, 3), (2, 0), (0, 10**6)):
                    lo, hi = min(r0 * 4, strips), min((r0 + cnt) * 4, strips)
                    want = [k for k, s_ in enumerate(mine) if lo <= s_ < hi]
                    first, n = band(h, r, world, r0, cnt)
                    assert n == len(want), (h, world, r, r0, cnt)
                    if n:
                        assert first == want[0]
            assert sorted(owned) == list(range(strips))
            assert lib.rtx_shard_strip_count(h, world, world) == 0  # index out of range
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=1)
    w, h = 40, 27  # ragged: 6 full strips + one of 3 rows
    img = oracle.render_scene(sc, w, h, seed=2, nthreads=2)["rgb"]
    strips = (h + 3) // 4
    for world in (2, 3):
        out = np.full_like(img, -1.0)
        flat_src, flat_dst = img.reshape(-1), out.reshape(-1)
        se = 4 * w * 3  # doubles per full strip
        for r in range(world):
            first, n = band(h, r, world, 0, 10**6)
            g0, g_last = first * world + r, (first + n - 1) * world + r
            ragged = g_last == strips - 1 and h % 4 != 0
            for j in range(n - 1 if ragged else n):  # the 2-D copy: row j at pitch world * se
                o = (g0 + j * world) * se
                flat_dst[o:o + se] = flat_src[o:o + se]
            if ragged:
                o = g_last * se
                flat_dst[o:] = flat_src[o:]
        assert out.tobytes() == img.tobytes(), world
