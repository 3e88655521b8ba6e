#!/usr/bin/env python
"""bench.py — the headline benchmark: Mrays/s on the sphere demo scene (S1) at 1920x1080, 64 spp, max
depth 8 (BASELINE.json config 2), on N B200s of one node.

  python bench.py --gpus 1 --steps K --warmup W                  (N=1)
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...                           (the reference's CPU path on host cores)

A "step" = one full render of the image.  "ray" = one nearest-hit query (closest_object, scene.rs:231),
counted on the device.  For N > 1 the image's 16x16 tiles are interleaved over the ranks (strong scaling:
the work per step is fixed), each rank renders its tiles with no data-path collective, then one NCCL
gather brings the shards to rank 0 and a de-interleave kernel assembles the image — all inside the step.
Rank 0 prints ONE JSON line.
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
# one `ncu --set full` capture of render_pool_kernel<64,1,true,24> on this workload (profiles/r01_render_pool_kernel_v6_raw.csv)
NCU_TRAFFIC = 457520000 + 674960000
NCU_TRAFFIC_SOURCE = ("bytes per launch, dram__bytes_read.sum 457.5 MB + dram__bytes_write.sum 675.0 MB of one ncu capture of "
                      "render_pool_kernel<64,1,true,24> on this workload (profiles/r01_SUMMARY.md): the sample queue's scratch is a "
                      "flag byte per camera path (132.7 MB) + 24 B of radiance for the 21 % of paths whose radiance is not zero; "
                      "tail_sum_kernel then reads 1.27 GB in 0.235 ms (5.4 TB/s) and writes the 49.8 MB frame")
METRIC = "Mrays/s at 1080p 64spp, 1/2/4/8 B200, vs ref CPU path on host cores"
WORKLOAD = "S1 sphere demo scene (1 plane + 6 spheres), 1920x1080, 64 spp, max depth 8 (BASELINE.json configs[1])"


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
def cpu_render_rate(width, height, target_seconds):
    """Times the CPU restatement of the reference path (oracle/, all host threads, pool of rows) on a
    bounded sample of the workload: the same scene/size/depth at a reduced spp.  Returns a dict."""
    from oracle import oracle
    from rtx_b200 import scenes
    cores = oracle.hardware_threads()
    sc = scenes.sphere_demo(rays_per_pixel=1, max_bounces=MAX_BOUNCES)
    t0 = time.perf_counter()
    r = oracle.render_scene(sc, width, height, seed=SEED)
    dt1 = time.perf_counter() - t0
    spp = int(max(1, min(SPP, target_seconds / max(dt1, 1e-3))))
    sc = scenes.sphere_demo(rays_per_pixel=spp, max_bounces=MAX_BOUNCES)
    t0 = time.perf_counter()
    r = oracle.render_scene(sc, width, height, seed=SEED)
    dt = time.perf_counter() - t0
    return {"value": r["rays"] / dt / 1e6, "unit": "Mrays/s", "cores": int(cores), "kind": "port",
            "sample": f"{width}x{height} x {spp} spp of {SPP} (same scene, seed, depth; {r['rays']} rays in {dt:.2f} s; "
                      f"C++ restatement of the Rust CPU path, thread pool over rows, no Mutex/dyn overhead)",
            "seconds": dt, "rays": int(r["rays"]), "spp": spp}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be
    compiled here (no cargo/rustc), so this is the oracle port; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import __graft_entry__ as g
    g.build()
    for _ in range(args.warmup):
        cpu_render_rate(args.width, args.height, 1.0)
    per_step = max(2.0, min(15.0, 120.0 / max(1, args.steps)))
    vals = [cpu_render_rate(args.width, args.height, per_step) for _ in range(args.steps)]
    rays = sum(v["rays"] for v in vals)
    secs = sum(v["seconds"] for v in vals)
    value = rays / secs / 1e6
    cb = dict(vals[-1])
    cb["value"] = value
    for k in ("seconds", "rays", "spp"):
        cb.pop(k)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "width": args.width, "height": args.height, "spp": args.spp,
                   "max_bounces": MAX_BOUNCES, "note": "each step is a bounded sample: same image at reduced spp"},
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

    if world == 1:
        rgb = torch.empty((H, W, 3), dtype=torch.float64, device=dev)
        ids = torch.empty((H, W), dtype=torch.int32, device=dev)
    elif args.gather == "nccl":
        rgb_c, ids_c = rdist.alloc_shard_buffers(W, H, world, dev, want_ids=True)
    else:
        frame = rdist.PeerFrame(W, H, local_rank, None, 0, want_ids=True)
        frame_rgb = frame.tensors()[0] if rank == 0 else None

    def step(seed):
        """One pass of the hot path over the whole image, outputs resident in HBM (on rank 0 for N > 1)."""
        if world == 1:
            scene.render_device(W, H, seed, rtx_b200.RTX_ACCEL_AUTO, None, rgb.data_ptr(), ids.data_ptr(),
                                counters.data_ptr(), stream.cuda_stream)
            return rgb
        if args.gather == "p2p":
            # gather fused into the render kernel: peers store finished pixels into rank 0's frame over NVLink
            frame.render(scene, seed, counters)
            return frame_rgb
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
        step(SEED)
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
        step(SEED)
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

    # ---- e2e: the public host-buffer API, pinned host memory, copies inside the timed region ----
    e2e = None
    npix = W * H
    pin = torch.empty((H, W, 3), dtype=torch.float64).pin_memory() if rank == 0 else None
    types, shapes, mats = scene.objects_flat()

    def e2e_step():
        if world == 1:
            # the step's inputs go host → device again: rebuild the scene's objects, then render into
            # pinned host memory through rtx_scene_render (D2H of the whole frame inside the call)
            lib = rtx_b200._abi.load()
            h = scene._sync()
            rtx_b200._abi.check(lib.rtx_scene_clear_objects(h), h)
            scene._synced_batches = 0
            scene.render(W, H, seed=SEED, out=pin.numpy())
            return scene.last_stats.rays
        if args.gather == "p2p":
            # the frame goes to rank 0's pinned host memory band by band while the next band renders on every rank
            frame.render_to_host(scene, SEED, pin, counters)
            return None
        out = step(SEED)
        if rank == 0:
            pin.copy_(out, non_blocking=True)
        return None

    for _ in range(min(WU, 2)):
        e2e_step()
    sync_all()
    t0 = time.perf_counter()
    for _ in range(K):
        e2e_step()
    sync_all()
    e2e_s = time.perf_counter() - t0
    # the frame the e2e call left in host memory is the frame the device-timed call left in HBM
    if world == 1:
        e2e_checked = bool(torch.equal(pin, rgb.cpu()))
    elif args.gather == "p2p" and rank == 0:
        e2e_checked = bool(torch.equal(pin, frame_rgb.cpu()))
    else:
        e2e_checked = None
    if e2e_checked is False:
        raise SystemExit("e2e frame (host buffer) differs from the device-resident frame")
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    rays_per_step = rays // K
    blob_bytes = (len(types) * (8 * 8) + int((types == 0).sum()) * 36 +
                                             int((types == 1).sum()) * 84 + int((types == 2).sum()) * 116)
    e2e = {"value": rays_per_step * K / float(te.item()) / 1e6, "unit": "Mrays/s",
           "h2d_bytes_per_step": int(blob_bytes), "d2h_bytes_per_step": int(npix * 24 + 32),
           "api": "rtx_scene_render (host buffers, pinned)" if world == 1 else
                  ("rtx_scene_render_device_band per rank (RTX_FLAG_SHARD_IN_PLACE into rank 0's peer frame) + banded D2H on rank 0"
                   if args.gather == "p2p" else
                   "rtx_scene_render_device per rank + NCCL gather + rtx_deinterleave_tiles + D2H on rank 0"),
           "ms_per_step": float(te.item()) / K * 1e3, "frame_equals_device_frame": e2e_checked,
           "d2h": ("inside rtx_scene_render: the frame is rendered in 4 bands of tile rows into a device frame; each band's "
                   "cudaMemcpyAsync to the pinned caller buffer runs on a copy stream while the next band renders"
                   if world == 1 else
                   ("PeerFrame.render_to_host: 4 bands; after a band's completion barrier rank 0's copy engine moves its rows to "
                    "pinned host memory while every rank renders the next band" if args.gather == "p2p" else
                    "cudaMemcpyAsync of rank 0's frame to pinned host memory after the step"))}

    if world > 1 and args.gather == "p2p":
        frame.close()
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (render_pool_kernel for S1), measured live ----
    nf_tf, fma_tf = rtx_b200.probe_fp64_peak(local_rank)
    flops, per_ray = algorithmic_flops(scene, rays, hits, paths)
    # at N > 1 dev_ms contains the gather; the kernel-only figure is quoted at N = 1
    achieved = flops / (dev_ms * 1e-3) / 1e12
    per_rank = max(1, scene.last_device_launches)
    sample_queue = per_rank >= 2  # (granule, sample) entries: 24 B of scratch per camera path, written then summed
    hbm_bytes = K * npix * 28 + (paths * 2 if sample_queue else 0)  # + the non-zero paths' 2 x 24 B (scene dependent)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    roofline = {
        "bound": "fp64", "achieved": achieved, "peak": nf_tf, "unit": "TFLOP/s", "frac": achieved / nf_tf if nf_tf else None,
        "traffic": NCU_TRAFFIC if (W, H, args.spp, world) == (WIDTH, HEIGHT, SPP, 1) and sample_queue else None,
        "traffic_source": NCU_TRAFFIC_SOURCE,
        "kernel": "rtx::render_pool_kernel (context-pool scheduler; S1 blob <= 16 KB, scanned from shared memory)" +
                  (" + rtx::tail_sum_kernel (ordered per-pixel sum of the sample queue's scratch, ~4 % of the step): the "
                   "events bracket both, so `achieved` is a lower bound for the render kernel alone" if sample_queue else ""),
        "launches": K * per_rank, "avg_launch_ms": dev_ms / K,
        "model": f"flops = rays*{per_ray} + 92*hits + 27*paths (FP64, non-fused; SURVEY.md 8d); div/sqrt/sincos count 1 each",
        "peak_source": "measured live: rtx_probe_fp64_peak non-fused DADD+DMUL issue rate (MEASURED_PEAKS.json has no FP64 figure)",
        "fp64_fma_peak_tflops": fma_tf,
        "note": "the path is FP64-CUDA-core bound, neither HBM nor tensor (SURVEY.md 8d); HBM view below for completeness",
        "hbm": {"achieved": hbm_bytes / (dev_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                "frac": hbm_bytes / (dev_ms * 1e-3) / 1e9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"},
    }
    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_baseline = cpu_render_rate(W, H, args.cpu_seconds)
        for k in ("seconds", "rays", "spp"):
            cpu_baseline.pop(k)
    # our kernels per step: what rtx_scene_render_device launched on a rank (render kernel; + the ordered-sum pass
    # under the sample queue) x ranks, + de-interleave on rank 0 with --gather nccl; NCCL's kernels are not ours
    per_rank = max(1, scene.last_device_launches)
    launches_per_step = world * per_rank if (world == 1 or args.gather == "p2p") else world * per_rank + 1
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": K, "warmup": WU,
        "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "width": W, "height": H, "spp": args.spp, "max_bounces": MAX_BOUNCES,
                   "seed": SEED, "parallelism": f"image tiles 16x16 interleaved over {world} GPU(s)" +
                   ("" if world == 1 else (", pixels stored into rank 0's frame by the render kernel over NVLink "
                                           "peer memory + 4-byte all-reduce as completion barrier, inside the step"
                                           if args.gather == "p2p" else ", NCCL gather to rank 0 inside the step")),
                   "l2": "flushed between timed iterations (256 MB memset outside the event pairs)",
                   "rays_per_step": rays_per_step, "hits_per_step": hits // K, "paths_per_step": paths // K},
        "e2e": e2e, "gpu_launches": launches_per_step * K, "roofline": roofline, "cpu_baseline": cpu_baseline,
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
