#!/usr/bin/env python
"""bench_preview.py — the f32 preview kernel (rtx_scene_render_preview_f32_device, the wgpu backend's arithmetic) on S1 at
1920x1080, 16 spp, max_bounces 8: device-timed Mrays/s (ray = one closest_object call of base_shader.wgsl:190), next to
the f64 path's figure for the same scene.  One JSON line.  Usage: python tools/bench_preview.py [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
import torch  # noqa: E402

from rtx_b200 import scenes  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda", 0)
W, H, SPP = 1920, 1080, 16
sc = scenes.sphere_demo(rays_per_pixel=SPP, max_bounces=8)
out = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
cnt = torch.zeros(8, dtype=torch.int64, device=dev)
st = torch.cuda.current_stream(dev)
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
for i in range(-2, steps):
    if i == 0:
        torch.cuda.synchronize(dev)
        cnt.zero_()
    if i >= 0:
        ev[i][0].record(st)
    sc.render_preview_device(W, H, out.data_ptr(), None, cnt.data_ptr(), st.cuda_stream)
    if i >= 0:
        ev[i][1].record(st)
torch.cuda.synchronize(dev)
ms = sum(a.elapsed_time(b) for a, b in ev) / steps
rays = int(cnt[0].item()) // steps
print(json.dumps({"kernel": "rtx::preview_kernel", "workload": f"S1, {W}x{H}, {SPP} spp, max_bounces 8 (f32, WGSL semantics)",
                  "ms_per_frame": ms, "rays_per_frame": rays, "mrays_s": rays / ms / 1e3,
                  "finite": bool(torch.isfinite(out).all().item())}))
