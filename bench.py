#!/usr/bin/env python
"""bench.py — headline benchmark of the render hot path (BASELINE.json metric: Mpaths/s & Mrays/s,
4K, 1024 spp, 1/2/4/8 B200, beside the reference CPU renderer).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c5|c4|c2|c1]

One "step" = one whole frame of the workload (default: config 5, 3840x2160 at 1024 camera samples per
pixel).  With N > 1 (launched by torch.distributed.run, one rank per GPU) the 1024 samples are sharded
across ranks (rank g renders global samples g, g+N, ...), the accumulation buffers are summed to rank 0
with one NCCL reduce inside the timed step and rank 0 resolves the image: strong scaling.

Prints ONE JSON line (rank 0).  `value` = Mrays/s with the scene resident in HBM, device-timed with
CUDA events; `e2e` = the same through the host-buffer C-ABI calls (scene upload + render + image
download inside the timed region); `roofline` = the wavefront stage kernels against the measured FP32
issue peak; `cpu_baseline` = the unmodified reference (oracle/_ref) on this box's host cores, bounded.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "path-trace_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (width, height, spp, description)
    "c5": (3840, 2160, 1024, "config 5: HDR-env-lit glossy scene with image textures, 3840x2160, 1024 spp (synthetic 3200x1600 lat-long env, synthetic textures)"),
    "c4": (3840, 2160, 512, "config 4: deep CSG stress, 64 transformed spheres/planes in nested Union/Difference, textured, 3840x2160, 512 spp"),
    "c2": (1920, 1080, 256, "config 2: diffuse sphere on a plane under a constant sky, 1920x1080, 256 spp"),
    "c1": (1920, 1080, 1024, "config 1: makeWorld() default scene (synthetic env in place of the missing Serpentine_Valley_3k.hdr), 1920x1080"),
}


def build_world(b, name):
    from ptb200 import scenes
    return {"c5": scenes.config5, "c4": scenes.config4, "c2": scenes.config2, "c1": scenes.config1}[name](b)


def flops_per_ray(desc):
    """Frozen algorithmic work model (DESIGN.md 'work model'): every f32
    add/mul/cmp/div/sqrt counts 1, an FMA-able pair 2.  Sphere test 22 (the miss cost — normals are
    only computed for the one boundary hit), plane 14, each TransformedObject level 33 per leaf, 4 per
    merge-loop iteration (3 iterations per operator), 4 per scanned span (2 spans), and a shaded vertex
    (hit point 6, refractStrength 31, reflect 22, scatter sample 60, accumulate 12, normal 27)."""
    total = 0.0
    for i in range(desc.n_prims):
        p = desc.prims[i]
        total += 22 if p.type == 0 else 14
        if p.chain >= 0:
            total += 33 * desc.chains[p.chain].count
    n_ops = desc.n_ops - desc.n_prims
    total += 4 * 3 * n_ops + 4 * 2
    total += 6 + 31 + 22 + 60 + 12 + 27
    return total


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop = index, [], False

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        reasons = []
        for k, name in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap")):
            if any(r[k].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "power_w_max": max(float(r[2]) for r in self.rows), "reasons": reasons, "samples": len(self.rows)}


def cpu_reference_rate(name, seconds, width, height):
    """The reference's own CPU renderer (unmodified sources in oracle/_ref, else the oracle port) on
    all host cores: the 192x108-style lattice of the workload's image plane, rows until `seconds`."""
    import common
    cores = os.cpu_count() or 1
    step = int(os.environ.get("PTB200_REF_LATTICE", "4"))  # rows of the lattice are rendered until `seconds` is up
    nx, ny = width // step, height // step
    flags = open("/proc/cpuinfo").read() if os.path.exists("/proc/cpuinfo") else ""
    kind, info = "port", None
    if common.have_ref():
        fast = os.path.exists(common.REF_FAST_LIB) and " avx2" in flags and " fma" in flags
        lib = common.ref(fast)
        rb = common.ref_builder(fast)
        world = build_world(rb, name)
        _, info = common.ref_render_mt(lib, world, width, height, step // 2, step // 2, step, nx, ny, 1, 16, 99, cores, seconds)
        kind = "reference"
        build = "oracle/_ref/libptref_fast.so (-O3 -march=x86-64-v3)" if fast else "oracle/_ref/libptref.so (-O3, no FMA)"
    else:
        from ptb200 import device
        from ptb200.builder import product_builder
        flat = device.Flat(build_world(product_builder(), name))
        _, info = common.orc_render_mt(flat, width, height, step // 2, step // 2, step, nx, ny, 1, 16, 99, cores, seconds)
        build = "oracle/libptoracle.so (CPU restatement)"
    secs = max(info["seconds"], 1e-9)
    return {
        "value": info["rays"] / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
        "mpaths_per_s": info["pixels"] / secs / 1e6, "rays_per_camera_sample": info["rays"] / max(info["pixels"], 1),
        "sample": "%s, every %dth pixel of the %dx%d image plane, 1 camera sample/px (each = the reference's ~10^4-ray branching estimate), %d px in %.1f s; %s"
                  % (name, step, width, height, info["pixels"], secs, build),
    }


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    width, height, spp, desc = WORKLOADS[args.workload]
    per_step = float(os.environ.get("PTB200_REF_SECONDS", "6"))
    for _ in range(args.warmup):
        cpu_reference_rate(args.workload, min(per_step, 1.0), width, height)
    t0 = time.time()
    rays = pixels = secs = 0.0
    last = None
    for _ in range(args.steps):
        last = cpu_reference_rate(args.workload, per_step, width, height)
        rays += last["value"] * 1e6
        pixels += last["mpaths_per_s"] * 1e6
    value = rays / args.steps / 1e6
    wall = time.time() - t0
    base = dict(last)
    base["value"] = value
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sample": base["sample"]}, "mpaths_per_s": pixels / args.steps / 1e6, "cpu_baseline": base,
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
    }))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from ptb200 import capi, device
    from ptb200.builder import product_builder

    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    width, height, spp, desc = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    if spp % world_size:
        raise SystemExit("spp must be divisible by the number of GPUs")
    lib = capi.product()
    flat = device.Flat(build_world(product_builder(), args.workload))
    scene = device.DeviceScene(flat)
    stream = torch.cuda.current_stream().cuda_stream
    accum = torch.zeros((height, width, 4), device="cuda")
    rgba = torch.empty_like(accum)
    host_image = torch.empty((height, width, 4), pin_memory=True)
    params = device.default_params(width, height, spp // world_size, seed=0xB200, sample_offset=rank, sample_stride=world_size, samples_per_pass=args.samples_per_pass)

    from ptb200 import multi

    def make_params(**kw):
        return device.default_params(width, height, seed=0xB200, samples_per_pass=args.samples_per_pass, **kw)

    def step(want_stats):
        accum.zero_()  # 132 MB write (> L2): also evicts the previous step's lines
        return multi.render_sharded(lambda p, a: scene.render_accumulate(p, a, stream, want_stats), lambda a, o: device.resolve(a, o, stream),
                                    make_params, spp, accum, rgba, dist if world_size > 1 else None)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(False)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step(False)
    ev1.record()
    barrier()
    sampler.stop = True
    ms = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
    if world_size > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = ms.item() / args.steps

    # one extra instrumented step for ray counts and per-stage device times (CUDA events inside the library)
    barrier()
    st = step(True)
    barrier()
    spec_ready, spec_compile_ms, spec_passes, spec_total = scene.specialized()  # did the scene-specialised trace kernels run?
    counts = torch.tensor([st.rays, st.paths, st.launches, st.queue_bytes], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    rays, paths, launches, queue_bytes = counts.tolist()

    # end-to-end through the host-buffer C ABI: scene upload + render + image download, every step
    h2d = sum(4 * 4 * flat.desc.images[i].width * flat.desc.images[i].height for i in range(flat.desc.n_images))
    h2d += 32 * flat.desc.n_prims + 96 * flat.desc.n_xforms + 4 * flat.desc.n_ops + 32 * flat.desc.n_materials + 8 * flat.desc.n_texops + 4 * flat.desc.n_texdata
    e2e_steps = max(1, min(args.steps, 2))
    host_np = host_image.numpy()

    e2e_parts = [0.0, 0.0, 0.0]  # scene create, render + download, destroy (host clock, rank 0's own)

    def e2e_step():
        t_a = time.perf_counter()
        sc2 = device.DeviceScene(flat)
        t_b = time.perf_counter()
        if world_size == 1:
            sc2.render(params, host_np)
        else:
            accum.zero_()
            sc2.render_accumulate(params, accum, stream, False)
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
            if rank == 0:
                device.resolve(accum, rgba, stream)
                host_image.copy_(rgba, non_blocking=True)
        torch.cuda.synchronize()
        t_c = time.perf_counter()
        sc2.close()
        t_d = time.perf_counter()
        for k, v in enumerate((t_b - t_a, t_c - t_b, t_d - t_c)):
            e2e_parts[k] += v * 1e3 / e2e_steps

    e2e_step()
    e2e_parts[:] = [0.0, 0.0, 0.0]
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / e2e_steps], device="cuda")
    if world_size > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)

    if rank != 0:
        if world_size > 1:
            dist.destroy_process_group()
        return
    fma_per_s, implied_mhz = device.fp32_peak()
    fpr = flops_per_ray(flat.desc)
    stage_ms = st.ms_primary + st.ms_bounce
    # roofline kernel = the primary TRACE kernel (tracePrimaryKernel): its launches are timed alone with
    # CUDA events on the render stream inside the library (PtbStats.ms_trace_primary)
    n_trace_launches = max(1, int(st.launches) // (2 + 2 * 16))
    trace_ms = st.ms_trace_primary
    achieved = st.rays_primary * fpr / (trace_ms * 1e-3) / 1e12 if trace_ms > 0 else 0.0
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
    except Exception:
        pass
    peak = 2 * fma_per_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    value = rays / (ms_per_step * 1e-3) / 1e6
    cpu = cpu_reference_rate(args.workload, float(os.environ.get("PTB200_REF_SECONDS", "15")), width, height) if world_size == 1 and not args.no_cpu else None
    out = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "width": width, "height": height, "spp": spp, "max_depth": 16, "samples_per_pass": args.samples_per_pass or "auto",
                   "parallelism": "spp-sharded x%d + NCCL sum-reduce" % world_size if world_size > 1 else "1 GPU",
                   "l2": "accumulation buffer (133 MB) is rewritten and the ray queues (GBs) stream through HBM every step: working set >> 126 MB L2"},
        "mpaths_per_s": paths / (ms_per_step * 1e-3) / 1e6, "rays_per_path": rays / paths,
        "e2e": {"value": rays / (e2e_ms.item() * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_ms.item(), "parts_ms": {"scene_create": e2e_parts[0], "render_and_download": e2e_parts[1], "destroy": e2e_parts[2]}, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": width * height * 16, "what": "ptb200_scene_create (uploads scene + images) + ptb200_render into a host buffer + destroy, per step"},
        "gpu_launches": int(launches / world_size) * args.steps + args.steps * (1 if rank == 0 else 0),
        "clocks": sampler.summary(),
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None, "traffic": traffic,
                     "kernel": ("ptbSpecTracePrimary (scene-specialised, NVRTC)" if spec_passes == spec_total and spec_total else "tracePrimaryKernel") + " (rank 0): %d launches, %.3f ms average, %.0f Mrays/s" % (n_trace_launches, trace_ms / n_trace_launches, st.rays_primary / (trace_ms * 1e-3) / 1e6 if trace_ms > 0 else 0.0),
                     "flops_per_ray_model": fpr, "whole_wavefront": {"achieved": st.rays * fpr / (stage_ms * 1e-3) / 1e12 if stage_ms > 0 else 0.0, "stage_ms": stage_ms},
                     "peak_source": "ptb200_fp32_peak FMA-chain microbenchmark on this GPU (MEASURED_PEAKS.json has no FP32 entry); %.2f T lane-FMA/s = 128 lanes x %d SMs x %.0f MHz" % (fma_per_s / 1e12, torch.cuda.get_device_properties(0).multi_processor_count, implied_mhz),
                     "hbm_queues": {"bound": "hbm", "achieved": queue_bytes / world_size / (stage_ms * 1e-3) / 1e9 if stage_ms > 0 else 0.0, "peak": hbm_peak, "unit": "GB/s",
                                    "frac": queue_bytes / world_size / (stage_ms * 1e-3) / 1e9 / hbm_peak if stage_ms > 0 else None, "bytes_per_ray": "16 B hit record + 176 B per queued ray (64 written, 64 + 32 read, 16 hit)"}},
        "trace_kernels": {"specialized_passes": spec_passes, "passes": spec_total, "nvrtc_compile_ms": spec_compile_ms if spec_ready else None,
                          "policy": os.environ.get("PTB_SPECIALIZE", "auto: programs <= 96 words, compiled in the background during warm-up")},
        "stage_ms": {"primary": st.ms_primary, "bounce": st.ms_bounce, "total": st.ms_total, "trace_primary": st.ms_trace_primary, "trace_bounce": st.ms_trace_bounce},
        "cpu_baseline": cpu,
    }
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(out) + "\n").encode())
    if world_size > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (parity/debug runs only)")
    ap.add_argument("--samples-per-pass", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
