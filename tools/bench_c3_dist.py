#!/usr/bin/env python
"""bench_c3_dist.py — config C3 across N GPUs: S2 (10k random spheres), 3840x2160, 16 spp, depth 8, LBVH, image
tiles interleaved over the ranks, every rank's render kernel storing its pixels into rank 0's frame over NVLink
peer memory (rtx_b200/dist.py PeerFrame), 4-byte all-reduce as completion barrier.  Launch:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_c3_dist.py
(N = 1: plain `python tools/bench_c3_dist.py`).  Rank 0 prints one JSON line: device-timed (CUDA events, max
over ranks) Mrays/s for the whole job, plus a byte-identity check of the N-rank image against the 1-rank image
at 960x540 (the full-size identity is what tests/test_dist_gpu.py proves at 2 GPUs)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rtx_b200 import dist as rdist  # noqa: E402
from rtx_b200 import scenes  # noqa: E402

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29533")
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
W, H, STEPS = 3840, 2160, int(os.environ.get("STEPS", "3"))
sc = scenes.random_spheres(device=local)  # S2: 10k spheres, 16 spp, depth 8
stream = torch.cuda.current_stream(dev)

# identity check at reduced size
small = rdist.PeerFrame(960, 540, local, None, 0, want_ids=True)
small.render(sc, 1, None)
torch.cuda.synchronize()
dist.barrier()
identical = None
if rank == 0:
    rgb, ids = small.tensors()
    full, ids_full = sc.render(960, 540, seed=1, want_ids=True)
    identical = bool(rgb.cpu().numpy().tobytes() == full.tobytes() and
                     (ids.cpu().numpy().view("uint32") == ids_full).all())
small.close()

frame = rdist.PeerFrame(W, H, local, None, 0, want_ids=True)
counters = torch.zeros(8, dtype=torch.int64, device=dev)
align = torch.zeros(1, dtype=torch.int32, device=dev)
for _ in range(2):
    frame.render(sc, 1, None)
torch.cuda.synchronize()
dist.barrier()
ms = 0.0
for _ in range(STEPS):
    dist.all_reduce(align)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    frame.render(sc, 1, counters)
    e1.record(stream)
    torch.cuda.synchronize()
    ms += e0.elapsed_time(e1)
t = torch.tensor([ms], dtype=torch.float64, device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
dist.all_reduce(counters)
if rank == 0:
    rays = int(counters[0].item()) // STEPS
    per = float(t.item()) / STEPS
    print(json.dumps({"config": "C3", "scene": "S2 10k spheres, LBVH", "width": W, "height": H, "spp": 16, "n_gpus": world,
                      "ms_per_frame": per, "rays_per_frame": rays, "mrays_s": rays / per / 1e3,
                      "nodes_per_ray": int(counters[4].item()) / STEPS / rays,
                      "n_rank_image_equals_1_rank_image_960x540": identical,
                      "assembly": "peer stores into rank 0's frame from the render kernel + 4-byte all-reduce"}))
frame.close()
dist.barrier()
dist.destroy_process_group()
