"""Image-tile sharding across GPUs: one process per GPU (torch.distributed), no data-path collective
while rendering.  ≙ the row threads + row-ordered join of Scene::render (raytracing/scene.rs:149-169).
The default assembly is PeerFrame (below): every rank's render kernel stores its finished pixels into rank
0's frame over NVLink peer memory, and a 4-byte all-reduce is the completion barrier.  The NCCL path
(render_distributed: compact shard buffers, one gather to rank 0, a de-interleave kernel) is kept as the A/B
baseline.

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


# ---------------------------------------------------------------------------------------------------
# Fused render + gather over NVLink peer stores
# ---------------------------------------------------------------------------------------------------
class _DevArray:
    """Raw device memory exposed to torch through __cuda_array_interface__ (no copy)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}


class PeerFrame:
    """The full frame [h][w][3] f64 (+ [h][w] u32 ids) lives in rank `dst`'s HBM; every other rank maps it
    (rtx_peer_export / rtx_peer_open — cudaIpc) and its render kernel stores finished pixels at their final
    position through NVLink while it renders (RTX_FLAG_SHARD_IN_PLACE).  ≙ the row-ordered join of
    scene.rs:167-169 without a gather step: the only collective left is the completion barrier.

    The frame is DOUBLE-BUFFERED: render k goes to buffer k & 1.  The barrier only tells the owner that the peers
    have finished; nothing tells the peers that the owner has finished READING, and a peer starts its next render
    as soon as its own barrier completes.  With two buffers the peers' render k+1 goes to the other buffer, and
    their render k+2 starts after barrier k+1, which the owner enqueues on its stream behind everything it did with
    frame k (its consumer kernels, or the wait on the copy stream that render_to_host ends with).  So: consume a
    frame on the owner's CURRENT stream (or make that stream wait for the consumer) before calling render again.
    All work is enqueued on torch's current stream of the device — NCCL's all-reduce runs there too."""

    def __init__(self, width, height, device, group=None, dst=0, want_ids=True):
        self.width, self.height, self.device, self.group, self.dst = width, height, int(device), group, dst
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.want_ids = want_ids
        lib = _abi.load()
        self.rgb_bytes = width * height * 24
        self.ids_off = (self.rgb_bytes + 255) & ~255
        self.buf_bytes = (self.ids_off + (width * height * 4 if want_ids else 0) + 255) & ~255
        total = 2 * self.buf_bytes
        self._cur = 0  # buffer of the most recent render
        self.owner = self.rank == dst
        ptr = C.c_void_p()
        handle = torch.zeros(_abi.RTX_PEER_HANDLE_BYTES, dtype=torch.uint8)
        if self.owner:
            _abi.check(lib.rtx_peer_alloc(self.device, total, C.byref(ptr)))
            buf = (C.c_uint8 * _abi.RTX_PEER_HANDLE_BYTES)()
            _abi.check(lib.rtx_peer_export(self.device, ptr, buf))
            handle = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
        if self.world > 1:
            backend = dist.get_backend(group)
            h = handle.to(torch.device("cuda", self.device)) if backend == "nccl" else handle
            dist.broadcast(h, src=dist.get_global_rank(group, dst) if group is not None else dst, group=group)
            handle = h.cpu()
        if not self.owner:
            raw = (C.c_uint8 * _abi.RTX_PEER_HANDLE_BYTES).from_buffer_copy(bytes(handle.numpy().tobytes()))
            _abi.check(lib.rtx_peer_open(self.device, raw, C.byref(ptr)))
        self.base = int(ptr.value)
        self._token = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", self.device))

    @property
    def ptr(self):
        """Device address of the buffer the most recent render went to."""
        return self.base + self._cur * self.buf_bytes

    @property
    def rgb_ptr(self):
        return self.ptr

    @property
    def ids_ptr(self):
        return self.ptr + self.ids_off if self.want_ids else None

    def poison(self):
        """Owner: fills BOTH buffers with NaN / 0xFFFFFFFE (stream-ordered) — a pixel no rank writes then shows in
        any comparison (tests, bench.py's N-rank == 1-rank check).  Call on every rank; ends with a barrier so that
        no peer store can precede the fill."""
        dev = torch.device("cuda", self.device)
        if self.owner:
            raw = torch.as_tensor(_DevArray(self.base, (2 * self.buf_bytes // 8,), "<i8"), device=dev)
            raw.fill_(-2)  # 0xFF..FE: a NaN as f64, 0xFFFFFFFE / 0xFFFFFFFF as ids
        if self.world > 1:
            dist.all_reduce(self._token, group=self.group)

    def tensors(self):
        """torch views of the most recently rendered frame (valid on the owner until the next-but-one render; on
        peers they alias remote memory — do not read)."""
        dev = torch.device("cuda", self.device)
        rgb = torch.as_tensor(_DevArray(self.ptr, (self.height, self.width, 3), "<f8"), device=dev)
        ids = torch.as_tensor(_DevArray(self.ptr + self.ids_off, (self.height, self.width), "<i4"),
                              device=dev) if self.want_ids else None
        return rgb, ids

    def render(self, scene, seed, counters=None, flags=RTX_ACCEL_AUTO):
        """Every rank renders its tiles straight into the owner's frame, then a stream-ordered barrier:
        when it completes on the owner's stream every peer's kernel has finished, and a finished kernel's
        stores are visible system-wide.  Enqueued on torch's current stream."""
        stream = torch.cuda.current_stream(torch.device("cuda", self.device)).cuda_stream
        self._cur ^= 1
        scene.render_device(self.width, self.height, seed, flags | _abi.RTX_FLAG_SHARD_IN_PLACE,
                            (self.rank, self.world), self.rgb_ptr, self.ids_ptr,
                            counters.data_ptr() if counters is not None else None, stream)
        if self.world > 1:
            dist.all_reduce(self._token, group=self.group)

    def band_rows(self, bands):
        """Tile-row ranges [(first, end), ...] of `bands` bands; the last one is half as tall as the others (its
        copy-out is the one nothing overlaps)."""
        tiles_y = (self.height + RTX_TILE - 1) // RTX_TILE
        bands = max(1, min(int(bands), tiles_y))
        den = 2 * bands - 1
        edges = [tiles_y * 2 * b // den for b in range(bands)] + [tiles_y]
        return [(edges[b], edges[b + 1]) for b in range(bands) if edges[b + 1] > edges[b]]

    def render_to_host(self, scene, seed, host_rgb, counters=None, flags=RTX_ACCEL_AUTO, bands=4):
        """render() with the frame handed over to pinned host memory on the owner: the image is rendered in bands of
        tile rows; after a band's completion barrier the owner's copy engine moves its rows to `host_rgb`
        ([h][w][3] f64, pinned; None on the other ranks) while every rank is already rendering the next band.
        Returns after the work is enqueued; synchronize the device before reading `host_rgb`."""
        dev = torch.device("cuda", self.device)
        cur = torch.cuda.current_stream(dev)
        if self.owner and getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        self._cur ^= 1
        frame_rgb = self.tensors()[0] if self.owner else None
        for (r0, r1) in self.band_rows(bands):
            scene.render_device_band(self.width, self.height, seed, flags | _abi.RTX_FLAG_SHARD_IN_PLACE,
                                     (self.rank, self.world), r0, r1 - r0, self.rgb_ptr, self.ids_ptr,
                                     counters.data_ptr() if counters is not None else None, cur.cuda_stream)
            if self.world > 1:
                dist.all_reduce(self._token, group=self.group)
            if self.owner:
                y0, y1 = r0 * RTX_TILE, min(self.height, r1 * RTX_TILE)
                ev = torch.cuda.Event()
                ev.record(cur)
                self._copy_stream.wait_event(ev)
                with torch.cuda.stream(self._copy_stream):
                    host_rgb[y0:y1].copy_(frame_rgb[y0:y1], non_blocking=True)
        if self.owner:
            cur.wait_stream(self._copy_stream)

    def close(self):
        lib = _abi.load()
        if self.base:
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                dist.barrier(group=self.group)
            if self.owner:
                _abi.check(lib.rtx_peer_free(self.device, C.c_void_p(self.base)))
            else:
                _abi.check(lib.rtx_peer_close(self.device, C.c_void_p(self.base)))
            self.base = 0


def render_distributed_p2p(scene, width, height, seed, group=None, flags=RTX_ACCEL_AUTO, want_ids=True, frame=None):
    """Multi-GPU render with the gather fused into the render kernel (PeerFrame).  Returns
    (rgb, ids, counters, frame) — rgb/ids are views of the owner's frame on rank 0, None elsewhere."""
    dev = torch.device("cuda", scene.device)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    if frame is None:
        frame = PeerFrame(width, height, scene.device, group, 0, want_ids)
    frame.render(scene, seed, counters, flags)
    if frame.owner:
        rgb, ids = frame.tensors()
        return rgb, ids, counters, frame
    return None, None, counters, frame
