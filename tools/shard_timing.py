#!/usr/bin/env python
"""shard_timing.py — one GPU renders shard r of N of the C2 workload (S1, 1920x1080, 64 spp, depth 8) and
reports the device time and ray count per shard: separates in-kernel effects (ramp, tail, imbalance between
shards) from communication in the multi-GPU numbers."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
import torch  # noqa: E402

import rtx_b200  # noqa: E402
from rtx_b200 import dist as rdist  # noqa: E402
from rtx_b200 import scenes  # noqa: E402

W, H, SPP = 1920, 1080, int(os.environ.get("SPP", "64"))
dev = torch.device("cuda", 0)
sc = scenes.sphere_demo(rays_per_pixel=SPP, max_bounces=8, device=0)
stream = torch.cuda.current_stream(dev)
for world in (1, 2, 4, 8, 16):
    rgb, ids = rdist.alloc_shard_buffers(W, H, world, dev, want_ids=True)
    rows = []
    for r in range(min(world, 8)):
        counters = torch.zeros(8, dtype=torch.int64, device=dev)
        for _ in range(2):
            rdist.render_shard(sc, W, H, 1, r, world, rgb, ids, None)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record(stream)
        for _ in range(reps):
            rdist.render_shard(sc, W, H, 1, r, world, rgb, ids, counters)
        e1.record(stream)
        torch.cuda.synchronize()
        rows.append((e0.elapsed_time(e1) / reps, int(counters[0].item()) // reps))
    ms = [a for a, _ in rows]
    rays = [b for _, b in rows]
    print(json.dumps({"world": world, "ms_min": min(ms), "ms_max": max(ms), "rays_min": min(rays), "rays_max": max(rays),
                      "ideal_ms_from_world1": None, "mrays_s_at_max": sum(rays) / len(rays) * world / max(ms) / 1e3}))
