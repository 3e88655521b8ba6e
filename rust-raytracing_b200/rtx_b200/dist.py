"""Image-tile sharding across GPUs: one process per GPU (torch.distributed), no data-path collective
while rendering, one gather of frame-buffer shards to rank 0 (NCCL over NVLink), then a de-interleave
kernel.  ≙ the row threads + row-ordered join of Scene::render (raytracing/scene.rs:149-169).

Ownership: the image is cut into RTX_TILE x RTX_TILE tiles numbered row-major; tile t belongs to rank
t % world (interleaved, so sky tiles and object tiles spread evenly).  Each rank writes its tiles
compactly — local tile j = global tile j*world + rank at [j][16][16][3] — so all ranks hold
equal-sized buffers and the gather is a plain concatenation.  torch is plumbing here (device memory,
streams, process group); pixels are only ever produced by librtx_cuda.so.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import RTX_ACCEL_AUTO, RTX_TILE, _abi, shard_tile_count


def shard_slots(width, height, world):
    """Compact-buffer slots per rank (= tiles owned by rank 0, the maximum over ranks)."""
    return int(shard_tile_count(width, height, 0, world))


def alloc_shard_buffers(width, height, world, device, want_ids=True):
    slots = shard_slots(width, height, world)
    rgb = torch.zeros((slots, RTX_TILE, RTX_TILE, 3), dtype=torch.float64, device=device)
    ids = torch.full((slots, RTX_TILE, RTX_TILE), -1, dtype=torch.int32, device=device) if want_ids else None
    return rgb, ids


def render_shard(scene, width, height, seed, rank, world, rgb, ids, counters, flags=RTX_ACCEL_AUTO, stream=None):
    """Asynchronously renders rank's tiles into its compact buffers on `stream` (default: torch's current)."""
    if stream is None:
        stream = torch.cuda.current_stream(rgb.device).cuda_stream
    scene.render_device(width, height, seed, flags, (rank, world), rgb.data_ptr(),
                        ids.data_ptr() if ids is not None else None,
                        counters.data_ptr() if counters is not None else None, stream)


def gather_shards(t, group=None, dst=0):
    """Gathers equal-sized per-rank tensors to `dst` → [world, *t.shape] (None on other ranks)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return t.unsqueeze(0)
    out = torch.empty((world, *t.shape), dtype=t.dtype, device=t.device) if rank == dst else None
    gather_list = list(out.unbind(0)) if rank == dst else None
    dist.gather(t, gather_list, dst=dst, group=group)
    return out


def deinterleave(width, height, world, gathered_rgb, gathered_ids=None):
    """Rank-major gathered tiles → full images [h][w][3] f64 / [h][w] i32, on the device (CUDA kernel)."""
    ref = gathered_rgb if gathered_rgb is not None else gathered_ids
    if not ref.is_cuda:
        raise RuntimeError("deinterleave runs on the GPU only (librtx_cuda.so); got a CPU tensor")
    dev = ref.device
    out_rgb = torch.empty((height, width, 3), dtype=torch.float64, device=dev) if gathered_rgb is not None else None
    out_ids = torch.empty((height, width), dtype=torch.int32, device=dev) if gathered_ids is not None else None
    stream = torch.cuda.current_stream(dev).cuda_stream
    _abi.check(_abi.load().rtx_deinterleave_tiles(
        width, height, world,
        gathered_rgb.data_ptr() if gathered_rgb is not None else None,
        out_rgb.data_ptr() if out_rgb is not None else None,
        gathered_ids.data_ptr() if gathered_ids is not None else None,
        out_ids.data_ptr() if out_ids is not None else None, stream))
    return out_rgb, out_ids


def render_distributed(scene, width, height, seed, group=None, flags=RTX_ACCEL_AUTO, want_ids=True):
    """Full multi-GPU render: every rank renders its tiles, rank 0 receives the image.
    Returns (rgb, ids, counters) on rank 0 — (None, None, counters) elsewhere.  counters = this rank's
    uint64[4] {rays, paths, triangle_panics, hits} device tensor."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device("cuda", scene.device)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    if world == 1:
        rgb = torch.empty((height, width, 3), dtype=torch.float64, device=dev)
        ids = torch.empty((height, width), dtype=torch.int32, device=dev) if want_ids else None
        scene.render_device(width, height, seed, flags, None, rgb.data_ptr(),
                            ids.data_ptr() if ids is not None else None, counters.data_ptr(),
                            torch.cuda.current_stream(dev).cuda_stream)
        return rgb, ids, counters
    rgb_c, ids_c = alloc_shard_buffers(width, height, world, dev, want_ids)
    render_shard(scene, width, height, seed, rank, world, rgb_c, ids_c, counters, flags)
    g_rgb = gather_shards(rgb_c, group)
    g_ids = gather_shards(ids_c, group) if want_ids else None
    if rank != 0:
        return None, None, counters
    rgb, ids = deinterleave(width, height, world, g_rgb, g_ids)
    return rgb, ids, counters


def render_all_shards_on_one_gpu(scene, width, height, seed, world, flags=RTX_ACCEL_AUTO):
    """One GPU plays every rank in turn (test helper for the byte-identity gate): returns
    (rgb, ids, total_rays) assembled exactly as rank 0 would after the gather."""
    dev = torch.device("cuda", scene.device)
    slots = shard_slots(width, height, world)
    g_rgb = torch.zeros((world, slots, RTX_TILE, RTX_TILE, 3), dtype=torch.float64, device=dev)
    g_ids = torch.full((world, slots, RTX_TILE, RTX_TILE), -1, dtype=torch.int32, device=dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    for r in range(world):
        render_shard(scene, width, height, seed, r, world, g_rgb[r], g_ids[r], counters, flags)
    rgb, ids = deinterleave(width, height, world, g_rgb, g_ids)
    torch.cuda.synchronize(dev)
    return rgb, ids, int(counters[0].item())
