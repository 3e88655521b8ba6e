"""Generates the committed golden fixtures from the CPU oracle (oracle/rtx_oracle.cpp).

The reference is Rust and cannot be run in this image, and it ships no golden vectors, so these files pin
the ORACLE (and, on the GPU box, the CUDA path) against drift — not against the reference itself
("parity unpinned").  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))

from oracle import oracle  # noqa: E402
from rtx_b200 import scenes  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=8)
r = oracle.render_scene(sc, 96, 64, seed=1)
np.savez_compressed(os.path.join(HERE, "sphere_demo_96x64_1spp_seed1.npz"), rgb=r["rgb"], ids=r["ids"],
                    rays=np.uint64(r["rays"]), rays_per_pixel=r["rays_per_pixel"])

# WyRand known answers for the restated fastrand 2.3.0 generator (first 8 u64 / f64 of three seeds)
seeds = np.array([0, 1, 0x0EF6F79ED30BA75A], dtype=np.uint64)
u = np.stack([oracle.wyrand_u64(int(s), 8)[0] for s in seeds])
f = np.stack([oracle.wyrand_f64(int(s), 8)[0] for s in seeds])
np.savez_compressed(os.path.join(HERE, "wyrand_kat.npz"), seeds=seeds, u64=u, f64=f)
# the f32 preview (wgsl_oracle.cpp): S1 at 96x64, 2 spp, max_bounces 5 — pins that restatement (and, on the GPU box, the
# preview kernel) against drift; like the fixtures above it is generated from OUR restatement, not from the shader
sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=5)
r = oracle.render_preview_scene(sc, 96, 64)
np.savez_compressed(os.path.join(HERE, "preview_sphere_demo_96x64_2spp.npz"), rgb=r["rgb"], ids=r["ids"], rays=np.uint64(r["rays"]))
# and its RNG (base_shader.wgsl:99-132): seed + 6 draws for three uv values
uv = np.array([[-0.25, 0.125], [0.0, 0.0], [0.4999, -0.4999]], dtype=np.float32)
probe = [oracle.wgsl_rng_probe(float(a), float(b), 6) for a, b in uv]
np.savez_compressed(os.path.join(HERE, "wgsl_pcg_kat.npz"), uv=uv, seed=np.array([p[0] for p in probe], dtype=np.uint32),
                    u32=np.stack([p[1] for p in probe]), f32=np.stack([p[2] for p in probe]))
print("golden fixtures written:", os.listdir(HERE))
