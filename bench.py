#!/usr/bin/env python
"""bench.py — headline benchmark of the `cuda` path-tracing pipeline (contract: task statement; metric: BASELINE.json).

Workload (configs[1], "C2"): TestScene-procedural on the cuda pipeline, 1920x1080, 256 spp, max depth 8, same seeds as
the reference CPU path tracer. One STEP = one full render of that image (W*H*spp paths).
  value   Mpaths/s, device-resident: scene, camera and accumulator already in HBM; CUDA events around the K steps.
  e2e     the same metric through the C ABI with HOST buffers: per step the scene tables are uploaded from host memory
          (spk_cuda_upload_scene), the image is rendered, tone-mapped and read back (spk_cuda_resolve) inside the timing.
  N > 1   one process per GPU (torchrun); the spp range is interleaved over ranks (first_sample = rank, stride = N), the
          accumulators are summed on rank 0 with ONE NCCL reduce, rank 0 resolves. Total work fixed -> "strong".
  roofline   render kernel, HBM bound: algorithmic bytes (SURVEY.md 8d / DESIGN.md) from the kernel's own counters.
  cpu_baseline / --impl reference: the reference's CPU path tracer (oracle/_ref, else the oracle port) on the host cores,
          bounded sample of the same workload. oracle/ is loaded ONLY on those two legs.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, HEIGHT, SPP, BOUNCE = 1920, 1080, 256, 8
WORKLOAD = "TestScene-procedural, cuda pipeline, 1920x1080, 256 spp, max depth 8 (BASELINE.json configs[1])"
SCENE = "testscene"
# --workload: the headline (c2) is what the driver runs; c4 / c5 are the scaling configs of BASELINE.json (configs[3],
# configs[4]); --spp bounds them (stated in config.workload) so a run stays within minutes
WORKLOADS = {
    "c2": ("testscene", 1920, 1080, 256, 8,
           "TestScene-procedural, cuda pipeline, 1920x1080, 256 spp, max depth 8 (BASELINE.json configs[1])"),
    "c4": ("city", 3840, 2160, 1024, 8,
           "city-10M (16 meshes x 62,500 tris, 160 instances = 10.0M instanced triangles), 3840x2160, 1024 spp, "
           "max depth 8, diffuse+glossy (BASELINE.json configs[3])"),
    "c5": ("room", 1920, 1080, 256, 32,
           "closed room, textured materials, emissive panel, 1920x1080, 256 spp, max depth 32 + Russian roulette "
           "(BASELINE.json configs[4])"),
}


def select_workload(args):
    global WIDTH, HEIGHT, SPP, BOUNCE, WORKLOAD, SCENE
    SCENE, WIDTH, HEIGHT, SPP, BOUNCE, WORKLOAD = WORKLOADS[args.workload]
    if args.spp:
        WORKLOAD += f" -- bounded to {args.spp} spp of {SPP} for this run"
        SPP = args.spp


def build_scene():
    from sparkle_b200 import scenes
    sb, camkw = getattr(scenes, SCENE)()
    return sb, camkw


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(key, **match):
    """dram bytes per launch of the dominant kernel from the committed ncu pass (profiles/r1_dram_traffic.json), or None
    when this run's launch is not the one that was captured"""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_dram_traffic.json")) as f:
            rec = json.load(f)[key]
    except Exception:
        return None
    return rec["dram_bytes"] if all(rec.get(k) == v for k, v in match.items()) else None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        busy = [x for x in sm if x > 0.5 * max(mx or [1])] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def scene_host_bytes(sb):
    n = 0
    for m in sb.meshes:
        n += m.vertices.nbytes + m.normals.nbytes + m.tangents.nbytes + m.indices.nbytes + (m.uvs.nbytes if m.uvs is not None else 0)
    for b in sb.blas:
        n += b.nodes.nbytes + b.prim_ids.nbytes
    if sb.tlas is not None:
        n += sb.tlas.nodes.nbytes + sb.tlas.prim_ids.nbytes
    n += sum(px.nbytes for px, _ in sb.textures)
    n += 112 * len(sb.prims) + 64 * len(sb.materials)
    return int(n)


def load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    for kind in ("reference", "port"):
        if O.available(kind):
            return O.Oracle(kind)
    raise RuntimeError("no oracle library: make -C oracle")


def cpu_render_rate(desc, cam, spp, first_sample=0):
    """reference CPU path tracer on the host cores; returns (Mpaths/s, seconds, cores, kind)"""
    orc = load_oracle()
    cores = os.cpu_count() or 1
    workers = max(1, min(64, cores) - 1)  # min(thread cvar 64, hw) - 1 reserved main thread (TaskDispatcher.cpp:19-24)
    t0 = time.perf_counter()
    orc.render(desc, cam, WIDTH, HEIGHT, first_sample, spp, max_bounce=BOUNCE, threads=workers)
    dt = time.perf_counter() - t0
    return WIDTH * HEIGHT * spp / dt / 1e6, dt, workers, orc.kind


def run_reference(args):
    """--impl reference: the reference's own CPU path tracer, bounded sample per step (rank 0 only under torchrun)"""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from sparkle_b200.scene import orbit_camera
    sb, camkw = build_scene()
    desc = sb.finalize()
    cam = orbit_camera(WIDTH, HEIGHT, **camkw)
    spp_step = 16 if SCENE == "testscene" else 2  # ~1.7 s of CPU work per step on 15 workers (testscene)
    for w in range(args.warmup):
        cpu_render_rate(desc, cam, spp_step, first_sample=w * spp_step)
    t0 = time.perf_counter()
    cores = kind = None
    for s in range(args.steps):
        _, _, cores, kind = cpu_render_rate(desc, cam, spp_step, first_sample=(args.warmup + s) * spp_step)
    dt = time.perf_counter() - t0
    paths = WIDTH * HEIGHT * spp_step * args.steps
    value = paths / dt / 1e6
    sample = f"{spp_step} spp of {SPP} per step at {WIDTH}x{HEIGHT} ({args.steps} steps), extrapolates linearly in spp"
    line = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": sample},
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores,
                             "kind": "reference" if kind == "reference" else "port", "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


class AccumView:
    """torch view of the context's device accumulator (for the NCCL reduce), via __cuda_array_interface__"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                         "strides": None}


def run_cuda(args):
    import torch
    import torch.distributed as dist
    from sparkle_b200 import _abi as A
    from sparkle_b200 import multi_gpu
    from sparkle_b200.device import CudaContext
    from sparkle_b200.scene import orbit_camera

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the cuda pipeline has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    first_sample, spp_rank, stride = multi_gpu.sample_slice(SPP, rank, world)

    sb, camkw = build_scene()
    desc = sb.finalize()
    cam = orbit_camera(WIDTH, HEIGHT, **camkw)
    ctx = CudaContext(WIDTH, HEIGHT, device=local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)  # torch.cuda.Event sees this stream
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ptr, nbytes = ctx.accum_ptr()
    accum_t = torch.as_tensor(AccumView(ptr, nbytes), device=f"cuda:{local}")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # megakernel or wavefront: measured on this scene (spk_cuda_choose_tracer), the same choice SPK_FLAG_AUTO_TRACER makes
    if args.tracer == "auto":
        use_wavefront, choose_ms = ctx.choose_tracer(BOUNCE)
    else:
        use_wavefront, choose_ms = args.tracer == "wavefront", None
    tracer_flag = A.FLAG_WAVEFRONT if use_wavefront else 0

    def device_step():
        ctx.reset_accumulation()
        ctx.render(first_sample=first_sample, sample_count=spp_rank, sample_stride=stride, max_bounce=BOUNCE,
                   flags=tracer_flag)
        multi_gpu.reduce_accumulator(accum_t, world, dst=0)  # ONE NCCL reduce over NVLink

    def timed(step_fn, steps, warmup):
        for _ in range(warmup):
            flush.fill_(1)
            step_fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for k in range(steps):
            flush.fill_(k)  # L2 flush between timed iterations, outside the events
            ev[k][0].record()
            step_fn()
            ev[k][1].record()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local}")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- kernel counters (one untimed instrumented pass on a 1/16-spp sample) ----
    ctx.reset_stats()
    ctx.reset_accumulation()
    cspp = max(1, spp_rank // 16)
    ctx.render(first_sample=first_sample, sample_count=cspp, sample_stride=stride, max_bounce=BOUNCE,
               flags=A.FLAG_COUNTERS | tracer_flag)
    cst = ctx.stats()
    ctx.reset_stats()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.stats()["kernel_launches"]
    ms_total = timed(device_step, args.steps, args.warmup)
    # every step launches the same kernels: the timed share of (warm-up + timed) launches
    launches = (ctx.stats()["kernel_launches"] - launches0) * args.steps // (args.steps + args.warmup)
    st = ctx.stats()
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: host scene tables in, tone-mapped image out, every step ----
    h2d = scene_host_bytes(sb)
    d2h = WIDTH * HEIGHT * 8

    def e2e_step():
        ctx.upload_scene(desc)
        ctx.set_camera(cam)
        ctx.reset_accumulation()
        ctx.render(first_sample=first_sample, sample_count=spp_rank, sample_stride=stride, max_bounce=BOUNCE,
                   flags=tracer_flag)  # the tracer was chosen once for this scene, as CudaRenderer does at init
        multi_gpu.reduce_accumulator(accum_t, world, dst=0)
        if rank == 0:
            ctx.resolve(A.RESOLVE_TONEMAP_F16, total_samples=SPP)

    e2e_steps = max(1, min(args.steps, 3))
    ms_e2e = timed(e2e_step, e2e_steps, 1)

    paths_step = WIDTH * HEIGHT * SPP
    value = paths_step * args.steps / (ms_total * 1e-3) / 1e6
    e2e_value = paths_step * e2e_steps / (ms_e2e * 1e-3) / 1e6
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the render kernel (this rank's launch) ----
    peak, peak_src = peaks()
    scale = spp_rank / cspp
    rays_launch = cst["rays"] * scale
    paths_launch = WIDTH * HEIGHT * spp_rank
    alg_bytes = (64 * cst["node_visits"] + 48 * cst["tri_tests"] + 128 * cst["inst_visits"] + 16 * cst["sphere_tests"]) * scale \
        + 2 * 16 * WIDTH * HEIGHT
    kernel_ms = ms_total / args.steps  # the render kernel is the step at N=1 (memset + 1 launch)
    traffic = None
    if world == 1 and not use_wavefront and args.workload == "c2":
        traffic = recorded_traffic("c2", form=ctx.tracer_choice()[1], width=WIDTH, height=HEIGHT, spp_per_launch=SPP)
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    rays_total = st["rays"] / max(1, args.steps + args.warmup) * world
    line = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "parallelism": f"spp-split x{world}", "l2": "flushed between steps (256 MB fill)",
                   "tracer": "wavefront" if use_wavefront else "megakernel",
                   "megakernel_form": ctx.tracer_choice()[1],
                   "tracer_choice_ms_8spp": {"megakernel": choose_ms[0], "wavefront": choose_ms[1]} if choose_ms else None,
                   "mrays_per_s": rays_total / (kernel_ms * 1e-3) / 1e6, "rays_per_path": rays_launch / paths_launch},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / e2e_steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src, "kernel": ("wf_extend_kernel<CULL=true> (+ wf_shade/regen/advance; achieved is over the whole step)"
                                if use_wavefront else {831: "render_phased_kernel<CULL=true, 8 warps x 3 CTAs/SM>",
                                      1162: "render_sorted_kernel<CULL=true, 16 warps x 2 CTAs/SM>",
                                      1084: "render_sorted_kernel<CULL=true, 8 warps x 4 CTAs/SM>"}.get(
                                         ctx.tracer_choice()[1], "render_phased_kernel")),
                     "alg_bytes_per_ray": alg_bytes / max(1.0, rays_launch),
                     "node_visits_per_ray": cst["node_visits"] / max(1, cst["rays"]),
                     "tri_tests_per_ray": cst["tri_tests"] / max(1, cst["rays"])},
    }
    if not args.no_cpu_baseline and world == 1:
        cpu_spp = 128 if SCENE == "testscene" else 4  # ~13 s on 15 workers (testscene)
        v, dt, cores, kind = cpu_render_rate(desc, cam, cpu_spp)
        line["cpu_baseline"] = {"value": v, "unit": "Mpaths/s", "cores": cores,
                                "kind": "reference" if kind == "reference" else "port",
                                "sample": f"{cpu_spp} spp of {SPP} at {WIDTH}x{HEIGHT} ({dt:.1f} s), linear in spp"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_trace(args):
    """--workload c3 (BASELINE.json configs[2]): soup-1M, 64 M incoherent rays, closest hit only. One step = the whole
    batch; N ranks take contiguous slices, no collective. value: device-resident rays and hits; e2e: host buffers
    through spk_cuda_trace_closest on a 16 M-ray slice of the batch (stated in config)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from sparkle_b200 import _abi as A
    from sparkle_b200 import multi_gpu, scenes
    from sparkle_b200.device import CudaContext

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the cuda pipeline has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    total = 64 << 20
    begin, end = multi_gpu.ray_slice(total, rank, world)
    n = end - begin
    sb, _ = scenes.soup(1_000_000)
    desc = sb.finalize()
    ctx = CudaContext(64, 64, device=local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.upload_scene(desc)
    g = torch.Generator(device="cuda").manual_seed(2 + rank)
    rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 3.0 - 1.5
    d = torch.randn((n, 3), generator=g, device="cuda")
    rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
    del d
    hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for k in range(steps):
            ev[k][0].record()
            fn()
            ev[k][1].record()
        barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ctx.reset_stats()
    cn = min(n, 4 << 20)
    ctx.trace_closest_device(rays.data_ptr(), cn, hits.data_ptr(), A.FLAG_COUNTERS)
    cst = ctx.stats()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total = timed(lambda: ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), 0), args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    en = min(n, (16 << 20) // world)
    h_rays = rays[:en].cpu().pin_memory()  # pinned host buffers, as the contract asks
    h_hits = torch.empty((en, 5), dtype=torch.float32).pin_memory()

    def e2e_step():
        A.check(ctx.lib.spk_cuda_trace_closest(ctx.handle, h_rays.data_ptr(), en, h_hits.data_ptr(), 0))

    e2e_steps = max(1, min(args.steps, 3))
    ms_e2e = timed(e2e_step, e2e_steps, 1)
    if rank == 0:
        peak, peak_src = peaks()
        nv, tt = cst["node_visits"] / cn, cst["tri_tests"] / cn
        bytes_ray = 32 + 20 + 64 * nv + 48 * tt + 128 * cst["inst_visits"] / cn
        kernel_ms = ms_total / args.steps
        achieved = bytes_ray * n / (kernel_ms * 1e-3) / 1e9
        line = {"metric": "Mrays/s", "value": total * args.steps / (ms_total * 1e-3) / 1e6, "unit": "Mrays/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": kernel_ms,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "soup-1M (1,000,000 triangles, one mesh), 64 Mi incoherent rays, closest hit only "
                                       "(BASELINE.json configs[2])",
                           "parallelism": f"ray-slice x{world}", "l2": "inputs larger than L2 (2.1 GB rays, 1.3 GB hits per step)",
                           "e2e_sample": f"{en * world} rays through host buffers"},
                "clocks": clocks,
                "e2e": {"value": en * world * e2e_steps / (ms_e2e * 1e-3) / 1e6, "unit": "Mrays/s",
                        "h2d_bytes_per_step": en * 32, "d2h_bytes_per_step": en * 20, "ms_per_step": ms_e2e / e2e_steps},
                "gpu_launches": args.steps,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": recorded_traffic("c3", rays_per_launch=n) if world == 1 else None,
                             "peak_source": peak_src,
                             "kernel": "trace_persistent_kernel<CULL=true, VOTE=6>", "alg_bytes_per_ray": bytes_ray,
                             "node_visits_per_ray": nv, "tri_tests_per_ray": tt}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + ["c3"])
    ap.add_argument("--spp", type=int, default=0, help="bound the workload's spp (stated in config.workload)")
    ap.add_argument("--tracer", default="auto", choices=["auto", "megakernel", "wavefront"])
    args = ap.parse_args()
    if args.workload == "c3":
        args.warmup = max(args.warmup, 3)
        if args.impl == "cuda":
            return run_trace(args)
        raise SystemExit("bench.py: --workload c3 has no reference arm here (tests compare it with the oracle)")
    select_workload(args)
    args.warmup = max(args.warmup, 3) if args.impl == "cuda" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_cuda(args)


if __name__ == "__main__":
    main()
