#!/usr/bin/env python
"""render_demo.py — writes the sphere demo scene (S1) as binary PPM files, headless: the f64 path through
rtx_scene_render_to_rgb8 (≙ Scene::render_to_image, scene.rs:173-178) and the f32 preview (the wgpu backend's arithmetic,
quantised the same way — note that the reference's shader mirrors the picture left-right relative to its CPU path:
base_shader.wgsl:71 crosses with (0,0,1), camera.rs:44 with (0,0,-1)).  Usage: python tools/render_demo.py [width height spp] -> demo_f64.ppm, demo_preview.ppm"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
from rtx_b200 import scenes  # noqa: E402


def write_ppm(path, img):
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(np.ascontiguousarray(img, dtype=np.uint8).tobytes())


if __name__ == "__main__":
    w, h, spp = (int(a) for a in (sys.argv[1:4] + ["800", "600", "64"][len(sys.argv) - 1:])[:3])
    sc = scenes.sphere_demo(rays_per_pixel=spp, max_bounces=8)
    write_ppm("demo_f64.ppm", sc.render_to_image(w, h, seed=1))  # already flipped: row 0 = top
    pv = sc.render_preview(w, h)[::-1]  # row 0 = bottom -> top
    write_ppm("demo_preview.ppm", np.clip(np.nan_to_num(pv) * 256.0, 0, 255).astype(np.uint8))
    print(f"wrote demo_f64.ppm and demo_preview.ppm ({w}x{h}, {spp} spp; {sc.last_stats.rays} preview rays)")
