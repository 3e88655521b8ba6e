#!/usr/bin/env python
"""bench.py — the headline benchmark: Mrays/s on the sphere demo scene (S1) at 1920x1080, 64 spp, max
depth 8 (BASELINE.json config 2), on N B200s of one node.

  python bench.py --gpus 1 --steps K --warmup W                  (N=1)
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...                           (the reference's CPU path on host cores)

A "step" = one full render of the image (step i uses seed SEED + i).  "ray" = one nearest-hit query
(closest_object, scene.rs:231), counted on the device.  For N > 1 the image's 16x16 tiles are interleaved over the
ranks (strong scaling: the work per step is fixed), each rank renders its tiles with no data-path collective and its
render kernel stores the finished pixels into rank 0's frame over NVLink peer memory; a 4-byte all-reduce is the
completion barrier — all inside the step (--gather nccl: compact shards + NCCL gather + de-interleave kernel, the
A/B baseline).  After the timed loops, outside them: the N-rank frame and the end-to-end host frame of a fresh seed
are compared byte for byte with the frame rank 0 renders alone ("nrank_equals_1rank"), and the other BASELINE.json
configs (C3, C4, C5) are measured ("other_configs").  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-raytracing_b200"))

WIDTH, HEIGHT, SPP, MAX_BOUNCES, SEED = 1920, 1080, 64, 8, 1
# one `ncu --set full` capture of render_pool_kernel<64,1,true,24,false> on this workload in round 2
# (profiles/r02_render_pool_kernel_raw.csv: dram__bytes_read.sum 458.63 MB, dram__bytes_write.sum 675.98 MB)
NCU_TRAFFIC = 458631936 + 675983616
NCU_TRAFFIC_SOURCE = ("bytes per launch, dram__bytes_read.sum 458.6 MB + dram__bytes_write.sum 676.0 MB of one ncu capture of "
                      "render_pool_kernel<64,1,true,24,false> on this workload (profiles/r02_render_pool_kernel_raw.csv): the sample "
                      "queue's scratch is a flag byte per camera path (132.7 MB) + 24 B of radiance for the 21 % of paths whose "
                      "radiance is not zero; the reads are sector fills of those partial writes; tail_sum_kernel then reads 1.27 GB "
                      "in 0.235 ms (5.4 TB/s) and writes the 49.8 MB frame.  Over the 10 ms launch that is 0.11 TB/s: DRAM is not "
                      "what bounds this kernel (DESIGN.md 4.2)")
METRIC = "Mrays/s at 1080p 64spp, 1/2/4/8 B200, vs ref CPU path on host cores"
WORKLOAD = "S1 sphere demo scene (1 plane + 6 spheres), 1920x1080, 64 spp, max depth 8 (BASELINE.json configs[1])"


def config_dict(args):
    """The workload, printed identically by both arms (--impl b200 / --impl reference)."""
    return {"workload": WORKLOAD, "width": args.width, "height": args.height, "spp": args.spp,
            "max_bounces": MAX_BOUNCES, "seed": f"{SEED} + step index",
            "parallelism": f"b200 arm: image tiles 16x16 interleaved over {args.gpus} GPU(s); reference arm: image rows over "
                           "the host threads",
            "l2": "b200 arm: flushed between timed iterations (256 MB memset outside the event pairs); reference arm: n/a"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"],
                    help="N>1: p2p = peers store pixels into rank 0's frame from inside the render kernel "
                         "(NVLink peer memory) + one tiny all-reduce as completion barrier; nccl = compact shard "
                         "buffers + dist.gather + de-interleave kernel")
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--spp", type=int, default=SPP)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-other-configs", action="store_true", help="skip C3/C4/C5 (other_configs)")
    return ap.parse_args()


# ---- clocks --------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.thread, self.gpu = [], None, None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max((float(r[3]) for r in self.rows if len(r) > 3 and r[3].replace(".", "").isdigit()),
                                   default=None),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- reference arm / cpu baseline: the oracle on the host cores -----------------------------------
def cpu_render_rate(width, height, target_seconds, seed=SEED, both_threadings=True):
    """Times the CPU restatement of the reference path (oracle/) on a bounded sample of the workload — the same
    scene/size/depth at a reduced spp when the full 64 would take longer than the target — with a pool of
    hardware_threads() workers over the rows and, when asked, also with the reference's own one-thread-per-row
    (scene.rs:149-151: thread::spawn per y, 1080 threads here).  `value` is the faster of the two."""
    from oracle import oracle
    from rtx_b200 import scenes
    cores = oracle.hardware_threads()
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=MAX_BOUNCES)
    t0 = time.perf_counter()
    r = oracle.render_scene(sc, width, height, seed=seed)
    dt1 = time.perf_counter() - t0
    share = target_seconds / (2 if both_threadings else 1)
    spp = int(max(1, min(SPP, share / max(dt1, 1e-3))))
    sc = scenes.sphere_demo(rays_per_pixel=spp, max_bounces=MAX_BOUNCES)
    t0 = time.perf_counter()
    r = oracle.render_scene(sc, width, height, seed=seed)
    dt_pool = time.perf_counter() - t0
    dt_rows = None
    if both_threadings:
        t0 = time.perf_counter()
        r2 = oracle.render_scene(sc, width, height, seed=seed, nthreads=-1)
        dt_rows = time.perf_counter() - t0
        assert r2["rays"] == r["rays"]
    dt = dt_pool if dt_rows is None else min(dt_pool, dt_rows)
    return {"value": r["rays"] / dt / 1e6, "unit": "Mrays/s", "cores": int(cores), "kind": "port",
            "threading": {"pool_of_hardware_threads_mrays_s": r["rays"] / dt_pool / 1e6,
                          "thread_per_row_mrays_s": (r["rays"] / dt_rows / 1e6) if dt_rows else None,
                          "value_is": "pool" if dt_rows is None or dt_pool <= dt_rows else "thread per row (scene.rs:149-151)"},
            "sample": f"{width}x{height} x {spp} spp of {SPP} (same scene, seed, depth; {r['rays']} rays in {dt:.2f} s; "
                      f"C++ restatement of the Rust CPU path, no Mutex/dyn overhead)",
            "seconds": dt, "rays": int(r["rays"]), "spp": spp}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be
    compiled here (no cargo/rustc), so this is the oracle port; rank 0 only.  Every step renders with the pool of
    hardware threads; the first step also with the reference's thread per row, and the line's value is the faster
    rate of the two."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import __graft_entry__ as g
    g.build()
    for _ in range(args.warmup):
        cpu_render_rate(args.width, args.height, 1.0, both_threadings=False)
    per_step = max(2.0, min(15.0, 120.0 / max(1, args.steps)))
    vals = [cpu_render_rate(args.width, args.height, per_step * (2 if i == 0 else 1), seed=SEED + i, both_threadings=(i == 0))
            for i in range(args.steps)]
    rays = sum(v["rays"] for v in vals)
    secs = sum(v["rays"] / (v["threading"]["pool_of_hardware_threads_mrays_s"] * 1e6) for v in vals)
    pool_rate = rays / secs / 1e6
    row_rate = vals[0]["threading"]["thread_per_row_mrays_s"]
    value = max(pool_rate, row_rate or 0.0)
    cb = dict(vals[0])
    cb["value"] = value
    cb["threading"] = {"pool_of_hardware_threads_mrays_s": pool_rate, "thread_per_row_mrays_s": row_rate,
                       "value_is": "pool (all steps)" if pool_rate >= (row_rate or 0.0) else
                                   "thread per row (scene.rs:149-151; first step, faster than the pool's rate over all steps)"}
    spps = sorted({v["spp"] for v in vals})
    for k in ("seconds", "rays", "spp"):
        cb.pop(k)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(args),
        "sample_note": f"each step renders the whole image at {spps} spp of {args.spp} (a bounded sample when below "
                       f"{args.spp}); Mrays/s does not depend on spp",
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ---- the B200 arm -----------------------------------------------------------------------------------
def algorithmic_flops(scene, rays, hits, paths):
    """SURVEY.md §8(d) / DESIGN.md: FP64 non-fused flops of the hoisted minimal form.
    flops/ray = 19*N_sphere + 13*N_plane + 40*N_tri + 4 ; + 92 per hit ; + 27 per camera sample."""
    types, _, _ = scene.objects_flat()
    n_sph, n_pl, n_tri = int((types == 0).sum()), int((types == 1).sum()), int((types == 2).sum())
    per_ray = 19 * n_sph + 13 * n_pl + 40 * n_tri + 4
    return rays * per_ray + 92 * hits + 27 * paths, per_ray


def timed_frames(scene, w, h, flags, steps, warmup, dev, ids_only=False, seed=1):
    """Device-timed frames of one scene on one GPU (other_configs): outputs resident in HBM, scene and LBVH built
    before timing, one CUDA-event pair per frame on the launching stream."""
    import torch
    import rtx_b200
    rgb = None if ids_only else torch.empty((h, w, 3), dtype=torch.float64, device=dev)
    ids = torch.empty((h, w), dtype=torch.int32, device=dev)
    cnt = torch.zeros(8, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream(dev)
    fl = flags | (rtx_b200.RTX_FLAG_NO_RADIANCE if ids_only else 0)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(-warmup, steps):
        if i == 0:
            torch.cuda.synchronize(dev)
            cnt.zero_()
        if i >= 0:
            ev[i][0].record(st)
        scene.render_device(w, h, seed, fl, None, rgb.data_ptr() if rgb is not None else None, ids.data_ptr(),
                            cnt.data_ptr(), st.cuda_stream)
        if i >= 0:
            ev[i][1].record(st)
    torch.cuda.synchronize(dev)
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    c = [int(v) // steps for v in cnt.tolist()]
    return dict(ms_per_frame=ms, rays_per_frame=c[0], mrays_s=c[0] / ms / 1e3, nodes_per_ray=c[4] / max(c[0], 1),
                prims_per_ray=c[5] / max(c[0], 1), launches_per_frame=scene.last_device_launches)


def graph_replay_us(scene, w, h, dev, frames=64, replays=20):
    """The launch-bound end of C5 without the host in the loop: `frames` hit-id renders captured once into a CUDA graph
    (every rtx_scene_render_device call is one kernel node — the work queue resets itself in the kernel) and replayed;
    device-timed microseconds per frame."""
    import torch
    import rtx_b200
    ids = torch.empty((h, w), dtype=torch.int32, device=dev)
    side = torch.cuda.Stream(device=dev)
    fl = rtx_b200.RTX_ACCEL_AUTO | rtx_b200.RTX_FLAG_NO_RADIANCE
    scene.render_device(w, h, 1, fl, None, None, ids.data_ptr(), None, side.cuda_stream)  # resident before capture
    torch.cuda.synchronize(dev)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(frames):
            scene.render_device(w, h, 1, fl, None, None, ids.data_ptr(), None, side.cuda_stream)
    g.replay()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.current_stream(dev)
    e0.record(st)
    for _ in range(replays):
        g.replay()
    e1.record(st)
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) * 1e3 / (frames * replays)


def other_configs(world, rank, dev, local_rank, hbm_peak, nf_tf):
    """BASELINE.json configs[2..4], measured after the headline and outside its timed region.  N = 1: C3, C4 and the
    two ends of the C5 sweep on this GPU.  N > 1: C3 over the N ranks (PeerFrame, as the headline)."""
    import torch
    import torch.distributed as dist
    import rtx_b200
    from rtx_b200 import dist as rdist
    from rtx_b200 import scenes
    out = {}
    if world == 1:
        sc = scenes.random_spheres(device=local_rank)  # S2: 10k spheres, 16 spp, depth 8
        r = timed_frames(sc, 3840, 2160, rtx_b200.RTX_ACCEL_LBVH, 3, 1, dev)
        floor = 14 * 64 + 32  # ceil(log2 1e4) 64-B nodes + one 32-B sphere
        touched = r["nodes_per_ray"] * 64 + r["prims_per_ray"] * 32
        r.update(workload="C3: S2 10k random spheres, 3840x2160, 16 spp, depth 8, LBVH", n_gpus=1,
                 bytes_per_ray_floor=floor, bytes_per_ray_touched=touched,
                 roofline={"bound": "hbm", "achieved": floor * r["mrays_s"] * 1e6 / 1e9, "peak": hbm_peak, "unit": "GB/s",
                           "frac": floor * r["mrays_s"] * 1e6 / 1e9 / hbm_peak})
        out["C3"] = r
        del sc
        sc = scenes.terrain(grid=709, rays_per_pixel=1, max_bounces=4, device=local_rank)  # S3: 1 005 128 triangles
        r = timed_frames(sc, 1920, 1080, rtx_b200.RTX_ACCEL_LBVH, 5, 2, dev)
        floor = 20 * 64 + 72  # ceil(log2 1e6) 64-B nodes + one 72-B f64 triangle
        touched = r["nodes_per_ray"] * 64 + r["prims_per_ray"] * 112
        r.update(workload="C4: S3 1M-triangle terrain, 1920x1080, 1 spp, primary + 4 bounces, LBVH", n_gpus=1,
                 bytes_per_ray_floor=floor, bytes_per_ray_touched=touched,
                 roofline={"bound": "hbm", "achieved": floor * r["mrays_s"] * 1e6 / 1e9, "peak": hbm_peak, "unit": "GB/s",
                           "frac": floor * r["mrays_s"] * 1e6 / 1e9 / hbm_peak})
        out["C4"] = r
        del sc
        sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=0, device=local_rank)
        for name, (w, h) in (("C5_256x256", (256, 256)), ("C5_7680x4320", (7680, 4320))):
            r = timed_frames(sc, w, h, rtx_b200.RTX_ACCEL_AUTO, 30, 5, dev, ids_only=True)
            # per ray: 131 FP64 flops of the 7-object scan + a 4-byte id written
            if w == 256:
                r["us_per_frame_graph_replay"] = graph_replay_us(sc, w, h, dev)
            r.update(workload=f"C5: S1 primary-only hit ids, {w}x{h}, 1 spp", n_gpus=1, us_per_frame=r["ms_per_frame"] * 1e3,
                     roofline={"bound": "fp64", "achieved": r["mrays_s"] * 1e6 * 131 / 1e12, "peak": nf_tf, "unit": "TFLOP/s",
                               "frac": (r["mrays_s"] * 1e6 * 131 / 1e12 / nf_tf) if nf_tf else None})
            out[name] = r
        return out
    sc = scenes.random_spheres(device=local_rank)
    W, H, steps = 3840, 2160, 3
    frame = rdist.PeerFrame(W, H, local_rank, None, 0, want_ids=True)
    stream = torch.cuda.current_stream(dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    align = torch.zeros(1, dtype=torch.int32, device=dev)
    frame.render(sc, 1, None, rtx_b200.RTX_ACCEL_LBVH)
    torch.cuda.synchronize(dev)
    dist.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        dist.all_reduce(align)
        ev[i][0].record(stream)
        frame.render(sc, 2 + i, counters, rtx_b200.RTX_ACCEL_LBVH)
        ev[i][1].record(stream)
    torch.cuda.synchronize(dev)
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(counters)
    frame.close()
    if rank == 0:
        rays = int(counters[0].item()) // steps
        per = float(t.item()) / steps
        floor = 14 * 64 + 32
        out["C3"] = dict(workload="C3: S2 10k random spheres, 3840x2160, 16 spp, depth 8, LBVH, tiles interleaved over "
                                  f"{world} GPUs, peer stores into rank 0's frame", n_gpus=world, ms_per_frame=per,
                         rays_per_frame=rays, mrays_s=rays / per / 1e3, nodes_per_ray=int(counters[4].item()) / steps / rays,
                         prims_per_ray=int(counters[5].item()) / steps / rays, bytes_per_ray_floor=floor,
                         roofline={"bound": "hbm", "achieved": floor * rays / per / 1e6, "peak": hbm_peak * world,
                                   "unit": "GB/s", "frac": floor * rays / per / 1e6 / (hbm_peak * world)})
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch N>1 with torch.distributed.run")
    # stdout carries ONE JSON line: whatever libraries write to file descriptor 1 meanwhile (NCCL's version banner at
    # communicator creation, build output) goes to stderr; the descriptor is restored for the final print
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if rank == 0:
        g.build()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the rtx CUDA path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL prints its version banner (NCCL_DEBUG=VERSION and every level above it) on STDOUT: send its log to
        # stderr so that stdout stays the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()

    import rtx_b200
    from rtx_b200 import dist as rdist
    from rtx_b200 import scenes
    W, H, K, WU = args.width, args.height, args.steps, max(args.warmup, 0)
    scene = scenes.sphere_demo(rays_per_pixel=args.spp, max_bounces=MAX_BOUNCES, device=local_rank)
    stream = torch.cuda.current_stream(dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    rgb = ids = None
    if world == 1 or rank == 0:  # rank 0 also renders alone for the N-rank == 1-rank check
        rgb = torch.empty((H, W, 3), dtype=torch.float64, device=dev)
        ids = torch.empty((H, W), dtype=torch.int32, device=dev)
    if world > 1 and args.gather == "nccl":
        rgb_c, ids_c = rdist.alloc_shard_buffers(W, H, world, dev, want_ids=True)
    elif world > 1:
        frame = rdist.PeerFrame(W, H, local_rank, None, 0, want_ids=True)

    def render_alone(seed):
        """The whole image on this GPU alone, device-resident."""
        scene.render_device(W, H, seed, rtx_b200.RTX_ACCEL_AUTO, None, rgb.data_ptr(), ids.data_ptr(), None,
                            stream.cuda_stream)
        return rgb

    def step(seed):
        """One pass of the hot path over the whole image, outputs resident in HBM (on rank 0 for N > 1)."""
        if world == 1:
            scene.render_device(W, H, seed, rtx_b200.RTX_ACCEL_AUTO, None, rgb.data_ptr(), ids.data_ptr(),
                                counters.data_ptr(), stream.cuda_stream)
            return rgb
        if args.gather == "p2p":
            # gather fused into the render kernel: peers store finished pixels into rank 0's frame over NVLink
            frame.render(scene, seed, counters)
            return frame.tensors()[0] if rank == 0 else None
        rdist.render_shard(scene, W, H, seed, rank, world, rgb_c, ids_c, counters)
        g_rgb = rdist.gather_shards(rgb_c)
        g_ids = rdist.gather_shards(ids_c)
        if rank == 0:
            return rdist.deinterleave(W, H, world, g_rgb, g_ids)[0]
        return None

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(WU):
        step(SEED + i)
    sync_all()
    counters.zero_()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    align = torch.zeros(1, dtype=torch.int32, device=dev)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    sync_all()
    t_wall0 = time.perf_counter()
    for i in range(K):
        flush_buf.zero_()  # L2 flush between timed iterations (outside the event pair)
        if world > 1:
            dist.all_reduce(align)  # ranks leave the flush together: skew between ranks is not render time
        ev[i][0].record(stream)
        step(SEED + i)
        ev[i][1].record(stream)
    sync_all()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    cnt = counters.clone()
    t = torch.tensor([dev_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms = float(t.item())
    rays, paths, panics, hits = (int(v) for v in cnt.tolist()[:4])
    value = rays / (dev_ms * 1e-3) / 1e6
    per_rank = max(1, scene.last_device_launches)  # kernels one step launches on a rank

    # ---- e2e: the public host-buffer API, pinned host memory, copies inside the timed region ----
    e2e = None
    npix = W * H
    host_frame = None
    if world > 1 and args.gather == "p2p":
        # ONE pinned host frame in shared memory, mapped by every rank: each rank's copy engine delivers its own rows
        # over its own PCIe link (rtx_scene_render_rows_to_host) — no rank-0 funnel
        names = [f"/rtx-bench-{os.getpid()}"]
        dist.broadcast_object_list(names, src=0)
        if rank == 0:
            host_frame = rtx_b200.HostFrame(names[0], W, H, create=True)
        dist.barrier()
        if rank != 0:
            host_frame = rtx_b200.HostFrame(names[0], W, H, create=False)
        pin = torch.from_numpy(host_frame.array) if rank == 0 else None
    else:
        pin = torch.empty((H, W, 3), dtype=torch.float64).pin_memory() if rank == 0 else None
    types, shapes, mats = scene.objects_flat()
    e2e_rays = [0]

    def e2e_step(seed):
        if world == 1:
            # the step's inputs go host → device again: rebuild the scene's objects, then render into
            # pinned host memory through rtx_scene_render (D2H of the whole frame inside the call)
            lib = rtx_b200._abi.load()
            h = scene._sync()
            rtx_b200._abi.check(lib.rtx_scene_clear_objects(h), h)
            scene._synced_batches = 0
            scene.render(W, H, seed=seed, out=pin.numpy())
            e2e_rays[0] += scene.last_stats.rays
            return
        if args.gather == "p2p":
            # every rank renders its rows (strips of 4, interleaved over the ranks) band by band and its copy engine puts
            # each band's rows at their place in the shared host frame while the next band renders; the 4-byte all-reduce
            # is the stream-ordered completion barrier: behind it on rank 0's stream the frame is whole
            scene.render_rows_to_host(W, H, seed, rtx_b200.RTX_ACCEL_AUTO, (rank, world), host_frame.ptr, None,
                                      counters.data_ptr(), stream.cuda_stream)
            dist.all_reduce(align)
            return
        out = step(seed)
        if rank == 0:
            pin.copy_(out, non_blocking=True)

    for i in range(min(WU, 2)):
        e2e_step(SEED + i)
    sync_all()
    e2e_rays[0] = 0
    counters.zero_()
    t0 = time.perf_counter()
    for i in range(K):
        e2e_step(SEED + i)
    sync_all()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    cnt_e = counters.clone()
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt_e, op=dist.ReduceOp.SUM)
    e2e_total_rays = e2e_rays[0] if world == 1 else int(cnt_e[0].item())
    if e2e_total_rays != rays:
        raise SystemExit(f"e2e loop traced {e2e_total_rays} rays, the device-timed loop {rays}: not the same work")

    # ---- checks, outside the timed regions: a seed no timed step used, poisoned buffers ----
    # N > 1: the frame the N ranks assemble in rank 0's HBM, and the frame the e2e path leaves in rank 0's pinned host
    # memory, against the frame rank 0 renders alone.  N = 1: the e2e host frame against the device-resident frame.
    check_seed = SEED + K + 1000
    nrank_equals_1rank = None
    if pin is not None:
        pin.fill_(-1.0)
    if world > 1 and args.gather == "p2p":
        frame.poison()
    dev_frame = step(check_seed)
    dev_frame = dev_frame.clone() if dev_frame is not None else None
    sync_all()
    if world > 1 and args.gather == "p2p":
        frame.poison()
    e2e_step(check_seed)
    sync_all()
    if rank == 0:
        alone = render_alone(check_seed)
        torch.cuda.synchronize(dev)
        e2e_checked = bool(torch.equal(pin, alone.cpu()))
        if world > 1:
            nrank_equals_1rank = bool(torch.equal(dev_frame, alone)) and e2e_checked
            if not nrank_equals_1rank:
                raise SystemExit(f"{world}-rank frame differs from the frame rank 0 renders alone "
                                 f"(device frame equal: {bool(torch.equal(dev_frame, alone))}, host frame equal: {e2e_checked})")
        elif not (e2e_checked and torch.equal(dev_frame, alone)):
            raise SystemExit("e2e frame (host buffer) differs from the device-resident frame")
    else:
        e2e_checked = None

    e2e_rgb8 = None
    if world == 1:
        # the same step through rtx_scene_render_to_rgb8 (scene.rs:173-178 on the device): 3 bytes per pixel come back
        img8 = None
        for i in range(-1, K):
            if i == 0:
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
            img8 = scene.render_to_image(W, H, seed=SEED + max(i, 0))
        dt8 = time.perf_counter() - t0
        e2e_rgb8 = {"value": rays / dt8 / 1e6, "unit": "Mrays/s", "ms_per_step": dt8 / K * 1e3, "d2h_bytes_per_step": int(npix * 3),
                    "api": "rtx_scene_render_to_rgb8 (render + flip/x256/saturate on the device, uint8 image to host)"}
    blob_bytes = (len(types) * (8 * 8) + int((types == 0).sum()) * 36 +
                                             int((types == 1).sum()) * 84 + int((types == 2).sum()) * 116)
    e2e = {"value": e2e_total_rays / float(te.item()) / 1e6, "unit": "Mrays/s",
           "h2d_bytes_per_step": int(blob_bytes), "d2h_bytes_per_step": int(npix * 24 + 32),
           "api": "rtx_scene_render (host buffers, pinned)" if world == 1 else
                  ("rtx_scene_render_rows_to_host per rank into one shared pinned host frame (rtx_host_frame_*): every GPU "
                   "delivers its own rows over its own PCIe link" if args.gather == "p2p" else
                   "rtx_scene_render_device per rank + NCCL gather + rtx_deinterleave_tiles + D2H on rank 0"),
           "ms_per_step": float(te.item()) / K * 1e3, "frame_equals_device_frame": e2e_checked,
           "d2h": ("inside rtx_scene_render: the frame is rendered in 4 bands of tile rows into a device frame; each band's "
                   "cudaMemcpyAsync to the pinned caller buffer runs on a copy stream while the next band renders"
                   if world == 1 else
                   ("inside rtx_scene_render_rows_to_host on every rank: the rank's rows are rendered in bands into its own device "
                    "frame; after each band one cudaMemcpy2DAsync moves the band's strips to the host frame on a copy stream "
                    "while the next band renders; a 4-byte all-reduce per step is the completion barrier" if args.gather == "p2p" else
                    "cudaMemcpyAsync of rank 0's frame to pinned host memory after the step"))}

    if world > 1 and args.gather == "p2p":
        frame.close()
        pin = None
        host_frame.close()

    # ---- the other BASELINE.json configs (outside every timed region of the headline) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    nf_tf, fma_tf = rtx_b200.probe_fp64_peak(local_rank)
    others = None
    if not args.no_other_configs:
        del flush_buf
        torch.cuda.empty_cache()
        others = other_configs(world, rank, dev, local_rank, hbm_peak, nf_tf)
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (render_pool_kernel for S1), measured live ----
    flops, per_ray = algorithmic_flops(scene, rays, hits, paths)
    # at N > 1 dev_ms contains the completion barrier; N GPUs' flops are set against N GPUs' peak
    achieved = flops / (dev_ms * 1e-3) / 1e12
    peak = nf_tf * world
    sample_queue = per_rank >= 2  # (granule, sample) entries: scratch per camera path, written then summed
    hbm_bytes = K * npix * 28 + (paths * 2 if sample_queue else 0)  # + the non-zero paths' 2 x 24 B (scene dependent)
    roofline = {
        "bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
        "traffic": NCU_TRAFFIC if (W, H, args.spp, world) == (WIDTH, HEIGHT, SPP, 1) and sample_queue else None,
        "traffic_source": NCU_TRAFFIC_SOURCE,
        "kernel": "rtx::render_pool_kernel (context-pool scheduler; S1 blob <= 16 KB, scanned from shared memory)" +
                  (" + rtx::tail_sum_kernel (ordered per-pixel sum of the sample queue's scratch): the "
                   "events bracket both, so `achieved` is a lower bound for the render kernel alone" if sample_queue else ""),
        "launches": K * per_rank * world, "avg_launch_ms": dev_ms / K,
        "model": f"flops = rays*{per_ray} + 92*hits + 27*paths (FP64, non-fused; SURVEY.md 8d); div/sqrt/sincos count 1 each",
        "peak_source": "measured live: rtx_probe_fp64_peak non-fused DADD+DMUL issue rate of one GPU (MEASURED_PEAKS.json has "
                       f"no FP64 figure) x {world} GPU(s)",
        "fp64_fma_peak_tflops": fma_tf,
        "note": "the path is FP64-CUDA-core bound, neither HBM nor tensor (SURVEY.md 8d); HBM view below for completeness",
        "hbm": {"achieved": hbm_bytes / (dev_ms * 1e-3) / 1e9, "peak": hbm_peak * world, "unit": "GB/s",
                "frac": hbm_bytes / (dev_ms * 1e-3) / 1e9 / (hbm_peak * world),
                "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"},
    }
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_baseline = cpu_render_rate(W, H, args.cpu_seconds)
        for k in ("seconds", "rays", "spp"):
            cpu_baseline.pop(k)
    # our kernels per step: what rtx_scene_render_device launched on a rank (render kernel; + the ordered-sum pass
    # under the sample queue) x ranks, + de-interleave on rank 0 with --gather nccl; NCCL's kernels are not ours
    launches_per_step = world * per_rank if (world == 1 or args.gather == "p2p") else world * per_rank + 1
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": WU,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": config_dict(args),
        "assembly": None if world == 1 else
                    ("pixels stored into rank 0's frame by the render kernel over NVLink peer memory + 4-byte all-reduce as "
                     "completion barrier, inside the step" if args.gather == "p2p" else "NCCL gather to rank 0 + de-interleave "
                     "kernel, inside the step"),
        "counts": {"rays_per_step": rays / K, "hits_per_step": hits / K, "paths_per_step": paths / K},
        "e2e": e2e, "e2e_rgb8": e2e_rgb8, "gpu_launches": launches_per_step * K, "roofline": roofline, "cpu_baseline": cpu_baseline,
        "nrank_equals_1rank": nrank_equals_1rank, "check_seed": check_seed,
        "other_configs": others,
        "clocks": clocks, "wall_s_timed_region": t_wall, "triangle_panics": panics,
    }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
