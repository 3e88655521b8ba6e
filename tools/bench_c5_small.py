#!/usr/bin/env python
"""bench_c5_small.py — the launch-bound end of config C5 (S1, primary-only hit ids, 1 spp) at small sizes: device-timed
microseconds per frame of a CUDA-graph replay of 64 captured frames, for the kernel the library picks and for the lane
kernel (RTX_FLAG_LANE_KERNEL).  One JSON line per (size, kernel)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
import torch  # noqa: E402

import rtx_b200  # noqa: E402
from rtx_b200 import scenes  # noqa: E402


def replay_us(scene, w, h, flags, dev, frames=64, replays=20):
    ids = torch.empty((h, w), dtype=torch.int32, device=dev)
    side = torch.cuda.Stream(device=dev)
    scene.render_device(w, h, 1, flags, None, None, ids.data_ptr(), None, side.cuda_stream)
    torch.cuda.synchronize(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(frames):
            scene.render_device(w, h, 1, flags, None, None, ids.data_ptr(), None, side.cuda_stream)
    g.replay()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream(dev)
    e0.record(st)
    for _ in range(replays):
        g.replay()
    e1.record(st)
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) * 1e3 / (frames * replays), ids.cpu().numpy()


def main():
    dev = torch.device("cuda", 0)
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=0)
    base = rtx_b200.RTX_ACCEL_AUTO | rtx_b200.RTX_FLAG_NO_RADIANCE
    for (w, h) in [(64, 64), (256, 256), (512, 512), (1024, 1024), (1920, 1080), (3840, 2160), (7680, 4320)]:
        ref = None
        for name, fl in (("auto", base), ("lane", base | rtx_b200.RTX_FLAG_LANE_KERNEL)):
            us, ids = replay_us(sc, w, h, fl, dev, frames=64 if w * h <= 1 << 20 else 8)
            same = True if ref is None else bool((ids == ref).all())
            ref = ids if ref is None else ref
            print(json.dumps({"config": "C5", "width": w, "height": h, "kernel": name, "us_per_frame_graph_replay": us,
                              "mrays_s": w * h / us, "ids_equal_first_kernel": same}), flush=True)


if __name__ == "__main__":
    main()
