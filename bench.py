#!/usr/bin/env python
"""bench.py — headline benchmark of the `cuda` path-tracing pipeline (contract: task statement; metric: BASELINE.json).

Headline workload (configs[1], "C2"): TestScene-procedural on the cuda pipeline, 1920x1080, 256 spp, max depth 8, same
seeds as the reference CPU path tracer. One STEP = one full render of that image (W*H*spp paths).
  value   Mpaths/s, device-resident: scene, camera and accumulator already in HBM; CUDA events around the K steps.
  e2e     the same metric through the C ABI with HOST buffers: per step the scene tables are uploaded from host memory
          (spk_cuda_upload_scene), the image is rendered, the accumulators of all GPUs are combined, tone-mapped and read
          back into a pinned host frame inside the timing; `parts_ms` breaks one such step down.
  N > 1   one process per GPU (torchrun); the spp range is interleaved over ranks (first_sample = rank, stride = N) and the
          data plane is the LIBRARY's: spk_cuda_group_resolve (csrc/spk_group.cu) - every GPU sums its rows of all
          accumulators over NVLink peer memory (CUDA IPC), tone-maps them and stores them at rank 0, one fused kernel per
          GPU. torch.distributed only carries the peer handles once and the barrier / max-over-ranks of the timing.
          `--reduce nccl` is the round-1 path (one ncclReduce to rank 0 + rank-0 resolve) kept for A/B. Before timing,
          rank 0 checks the combined frame against a single-context render of the same samples.
  extra   the other configs BASELINE.json names, in the same JSON line at every N: c3 (soup-1M, 64 Mi incoherent rays) and
          its coherent counterpart (terrain-1M, 64 Mi camera rays), c4 (city-10M at 4K, bounded spp), c5 (room, depth 32).
  gpu_launches   this repo's kernels launched inside the timed steps (1 render kernel per step, +1 group kernel at N > 1;
          cub's radix-sort kernels that order the next launch's strips are library kernels and are not counted).
  roofline   dominant kernel, HBM bound by contract: algorithmic bytes (SURVEY.md 8d / DESIGN.md) from the kernel's own
          counters. `traffic` is the measured DRAM traffic of the same launch from the committed ncu pass when the launch
          matches it; `dram_frac` is that traffic over time over peak - the honest HBM utilisation figure.
  cpu_baseline / --impl reference: the reference's CPU path tracer (oracle/_ref, else the oracle port) on the host cores,
          bounded sample of the same workload. oracle/ is loaded ONLY on those two legs.
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

# scene, width, height, spp, max bounce, default bound on spp when run as an extra, description
WORKLOADS = {
    "c2": ("testscene", 1920, 1080, 256, 8, 0,
           "TestScene-procedural, cuda pipeline, 1920x1080, 256 spp, max depth 8 (BASELINE.json configs[1])"),
    "c4": ("city", 3840, 2160, 1024, 8, 64,
           "city-10M (16 meshes x 62,500 tris, 160 instances = 10.0M instanced triangles), 3840x2160, 1024 spp, "
           "max depth 8, diffuse+glossy (BASELINE.json configs[3])"),
    "c5": ("room", 1920, 1080, 256, 32, 0,
           "closed room, textured materials, emissive panel, 1920x1080, 256 spp, max depth 32 + Russian roulette "
           "(BASELINE.json configs[4])"),
}
KERNEL_NAMES = {831: "render_phased_kernel<CULL=true, 8 warps x 3 CTAs/SM>",
                1162: "render_sorted_kernel<CULL=true, 16 warps x 2 CTAs/SM, pixels owned by the CTA>",
                1084: "render_sorted_kernel<CULL=true, 8 warps x 4 CTAs/SM, pixels owned by the CTA>",
                2162: "render_sorted_kernel<CULL=true, 16 warps x 2 CTAs/SM, pixel-pulling>",
                2084: "render_sorted_kernel<CULL=true, 8 warps x 4 CTAs/SM, pixel-pulling>",
                3162: "render_sorted_kernel<CULL=true, 16 warps x 2 CTAs/SM, stable sort>",
                3084: "render_sorted_kernel<CULL=true, 8 warps x 4 CTAs/SM, stable sort>"}


class Env:
    """one process per GPU: ranks, device, torch handles, barrier"""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the cuda pipeline has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = f"cuda:{self.local}"
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, step_fn, steps, warmup, flush=True):
        """W untimed steps, then K steps between CUDA events on the launching stream; barrier + synchronize on both
        sides; the L2 is flushed (256 MB fill) before every step, outside the events; max over ranks"""
        torch = self.torch
        for _ in range(warmup):
            if flush:
                self.flush.fill_(1)
            step_fn()
        self.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for k in range(steps):
            if flush:
                self.flush.fill_(k)
            ev[k][0].record()
            step_fn()
            ev[k][1].record()
        self.barrier()
        t = torch.tensor([sum(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(key, **match):
    """dram bytes per launch of the dominant kernel from the committed ncu pass (profiles/r2_dram_traffic.json, made by
    tools/r2_dram_traffic.sh over this very command), or None when this run's launch is not the one that was captured"""
    for name in ("r2_dram_traffic.json", "r1_dram_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                rec = json.load(f)[key]
        except Exception:
            continue
        if all(rec.get(k) == v for k, v in match.items()):
            return rec["dram_bytes"]
    return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index, enabled=True):
        self.gpu, self.rows, self.proc, self.enabled = gpu_index, [], None, enabled

    def start(self):
        if not self.enabled:
            return self
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.enabled:
            return None
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


def cpu_render_rate(desc, cam, width, height, bounce, spp, first_sample=0):
    """reference CPU path tracer on the host cores; returns (Mpaths/s, seconds, cores, kind)"""
    orc = load_oracle()
    cores = os.cpu_count() or 1
    workers = max(1, min(64, cores) - 1)  # min(thread cvar 64, hw) - 1 reserved main thread (TaskDispatcher.cpp:19-24)
    t0 = time.perf_counter()
    orc.render(desc, cam, width, height, first_sample, spp, max_bounce=bounce, threads=workers)
    dt = time.perf_counter() - t0
    return width * height * spp / dt / 1e6, dt, workers, orc.kind


def run_reference(args):
    """--impl reference: the reference's own CPU path tracer, bounded sample per step (rank 0 only under torchrun)"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from sparkle_b200 import scenes
    from sparkle_b200.scene import orbit_camera
    scene, width, height, spp, bounce, _, workload = WORKLOADS[args.workload]
    sb, camkw = getattr(scenes, scene)()
    desc = sb.finalize()
    cam = orbit_camera(width, height, **camkw)
    spp_step = 16 if scene == "testscene" else 2  # ~1.7 s of CPU work per step on 15 workers (testscene)
    for w in range(args.warmup):
        cpu_render_rate(desc, cam, width, height, bounce, spp_step, first_sample=w * spp_step)
    t0 = time.perf_counter()
    cores = kind = None
    for s in range(args.steps):
        _, _, cores, kind = cpu_render_rate(desc, cam, width, height, bounce, spp_step,
                                            first_sample=(args.warmup + s) * spp_step)
    dt = time.perf_counter() - t0
    value = width * height * spp_step * args.steps / dt / 1e6
    sample = f"{spp_step} spp of {spp} per step at {width}x{height} ({args.steps} steps), extrapolates linearly in spp"
    line = {"impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload, "sample": sample,
                       "bvh": "this repo's binned-SAH tree (max_leaf 4) injected into the reference's traversal; "
                              "bvh::v2's own builder is absent from the mount (BASELINE.md)"},
            "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores,
                             "kind": "reference" if kind == "reference" else "port", "sample": sample},
            "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


class AccumView:
    """torch view of the context's device accumulator (for the NCCL A/B path), via __cuda_array_interface__"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                         "strides": None}


def measure_render(env, key, args, spp_bound=0, steps=None, warmup=None, cpu_baseline=False):
    """one path-tracing workload at env.world GPUs -> the fields of a bench line"""
    import numpy as np
    from sparkle_b200 import _abi as A
    from sparkle_b200 import multi_gpu, scenes
    from sparkle_b200.device import CudaContext
    from sparkle_b200.scene import orbit_camera

    torch, rank, world, local = env.torch, env.rank, env.world, env.local
    scene, width, height, spp, bounce, _, workload = WORKLOADS[key]
    if spp_bound and spp_bound < spp:
        workload += f" -- bounded to {spp_bound} spp of {spp} for this run"
        spp = spp_bound
    steps = steps or args.steps
    warmup = max(3, warmup if warmup is not None else args.warmup)
    first_sample, spp_rank, stride = multi_gpu.sample_slice(spp, rank, world)

    sb, camkw = getattr(scenes, scene)()
    desc = sb.finalize()
    cam = orbit_camera(width, height, **camkw)
    ctx = CudaContext(width, height, device=local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)  # torch.cuda.Event sees this stream
    ctx.upload_scene(desc)
    ctx.set_camera(cam)

    # megakernel or wavefront: measured on this scene (spk_cuda_choose_tracer), the same choice SPK_FLAG_AUTO_TRACER makes
    if args.tracer == "auto":
        use_wavefront, choose_ms = ctx.choose_tracer(bounce)
    else:
        use_wavefront, choose_ms = args.tracer == "wavefront", None
    tracer_flag = A.FLAG_WAVEFRONT if use_wavefront else 0

    # ---- multi-GPU data plane: the library's fused reduce + resolve over peer memory, or (A/B) NCCL from here ----
    group, reduce_kind, accum_t = None, "none (1 GPU)", None
    if world > 1:
        if args.reduce == "group":
            try:
                group = multi_gpu.ipc_group(ctx, rank, world)
                group.set_async(True)
                reduce_kind = "spk_cuda_group_resolve: fused reduce-scatter + tone map + gather over NVLink peer memory (CUDA IPC)"
            except A.SpkError as e:  # no peer access between these GPUs: say so and take the NCCL path
                reduce_kind = f"ncclReduce to rank 0 (group unavailable: {e})"
            ok = torch.tensor([1 if group else 0], device=env.dev)
            env.dist.all_reduce(ok, op=env.dist.ReduceOp.MIN)
            if not int(ok.item()) and group:
                group.close()
                group = None
                reduce_kind = "ncclReduce to rank 0 (group unavailable on another rank)"
        else:
            reduce_kind = "ncclReduce to rank 0 + rank-0 resolve (round-1 path, --reduce nccl)"
        if group is None:
            ptr, nbytes = ctx.accum_ptr()
            accum_t = torch.as_tensor(AccumView(ptr, nbytes), device=env.dev)
    frame = ctx.pinned_frame(A.RESOLVE_TONEMAP_F16) if rank == 0 else None

    def combine(readback):
        """all ranks' samples -> one tone-mapped frame (at rank 0; in `frame` when readback)"""
        if group is not None:
            group.resolve(A.RESOLVE_TONEMAP_F16, spp, out=frame, readback=readback)
        elif world > 1:
            ctx.accum_ptr()  # a pending reset must be real before NCCL reads the buffer
            multi_gpu.reduce_accumulator(accum_t, world, dst=0)
            if rank == 0 and readback:
                ctx.resolve(A.RESOLVE_TONEMAP_F16, total_samples=spp, out=frame)
        elif readback:
            ctx.resolve(A.RESOLVE_TONEMAP_F16, total_samples=spp, out=frame)

    # ---- N > 1: the combined frame is checked against one context holding the same samples, before any timing ----
    verify = None
    if world > 1:
        vs = 2 * world
        f, c, s = multi_gpu.sample_slice(vs, rank, world)
        ctx.reset_accumulation()
        ctx.render(first_sample=f, sample_count=c, sample_stride=s, max_bounce=bounce, flags=tracer_flag)
        if group is not None:
            lin = group.resolve(A.RESOLVE_LINEAR_F32, vs)
        else:
            ctx.accum_ptr()
            multi_gpu.reduce_accumulator(accum_t, world, dst=0)
            lin = ctx.resolve(A.RESOLVE_LINEAR_F32, total_samples=vs) if rank == 0 else None
        env.barrier()
        if rank == 0:
            ctx.reset_accumulation()
            ctx.render(first_sample=0, sample_count=vs, max_bounce=bounce, flags=tracer_flag)
            single = ctx.resolve(A.RESOLVE_LINEAR_F32)
            err = float(np.abs(lin - single).max())
            tol = 1e-6 * max(1.0, float(np.abs(single).max()))
            verify = {"samples": vs, "max_abs_diff_vs_one_gpu": err, "ok": bool(err <= 16 * tol)}
            if not verify["ok"]:
                raise SystemExit(f"bench.py: the {world}-GPU frame differs from the 1-GPU frame by {err}")
        env.barrier()

    def device_step():
        ctx.reset_accumulation()
        ctx.render(first_sample=first_sample, sample_count=spp_rank, sample_stride=stride, max_bounce=bounce,
                   flags=tracer_flag)
        combine(readback=False)

    # ---- kernel counters (one untimed instrumented pass on a 1/16-spp sample) ----
    ctx.reset_stats()
    ctx.reset_accumulation()
    cspp = max(1, spp_rank // 16)
    ctx.render(first_sample=first_sample, sample_count=cspp, sample_stride=stride, max_bounce=bounce,
               flags=A.FLAG_COUNTERS | tracer_flag)
    cst = ctx.stats()
    ctx.reset_stats()

    sampler = ClockSampler(local, enabled=rank == 0).start()
    launches0 = ctx.stats()["kernel_launches"]
    ms_total = env.timed(device_step, steps, warmup)
    # every step launches the same kernels: the timed share of (warm-up + timed) launches
    launches = (ctx.stats()["kernel_launches"] - launches0) * steps // (steps + warmup)
    st = ctx.stats()
    clocks = sampler.stop()

    # ---- e2e: host scene tables in, tone-mapped frame out (pinned host memory), every step ----
    h2d = scene_host_bytes(sb)
    d2h = width * height * 8

    def e2e_step():
        ctx.upload_scene(desc)
        ctx.set_camera(cam)
        ctx.reset_accumulation()
        ctx.render(first_sample=first_sample, sample_count=spp_rank, sample_stride=stride, max_bounce=bounce,
                   flags=tracer_flag)  # the tracer was chosen once for this scene, as CudaRenderer does at init
        combine(readback=True)

    e2e_steps = max(1, min(steps, 3))
    ms_e2e = env.timed(e2e_step, e2e_steps, 1)
    # one more step taken apart on the host clock (every part ends in a synchronize)
    parts = {}
    env.barrier()
    t0 = time.perf_counter()
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    parts["upload_scene"] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    ctx.reset_accumulation()
    ctx.render(first_sample=first_sample, sample_count=spp_rank, sample_stride=stride, max_bounce=bounce, flags=tracer_flag)
    ctx.synchronize()
    parts["render"] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    combine(readback=True)
    ctx.synchronize()
    parts["combine_tonemap_readback"] = (time.perf_counter() - t0) * 1e3
    env.barrier()

    paths_step = width * height * spp
    out = None
    if rank == 0:
        peak, peak_src = peaks()
        scale = spp_rank / cspp
        rays_launch = cst["rays"] * scale
        paths_launch = width * height * spp_rank
        alg_bytes = (64 * cst["node_visits"] + 48 * cst["tri_tests"] + 128 * cst["inst_visits"] + 16 * cst["sphere_tests"]) * scale \
            + 2 * 16 * width * height
        kernel_ms = ms_total / steps  # the render kernel is the step at N=1 (1 launch; the accumulator reset is free)
        form = ctx.tracer_choice()[1]
        traffic = None
        if world == 1 and not use_wavefront:
            traffic = recorded_traffic(key, form=form, width=width, height=height, spp_per_launch=spp)
        achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
        rays_total = st["rays"] / max(1, steps + warmup) * world
        out = {
            "value": paths_step * steps / (ms_total * 1e-3) / 1e6, "unit": "Mpaths/s", "ms_per_step": ms_total / steps,
            "steps": steps, "warmup": warmup,
            "config": {"workload": workload, "parallelism": f"spp-split x{world}", "reduce": reduce_kind,
                       "l2": "flushed between steps (256 MB fill)",
                       "tracer": "wavefront" if use_wavefront else "megakernel", "megakernel_form": form,
                       "tracer_choice_ms_8spp": {"megakernel": choose_ms[0], "wavefront": choose_ms[1]} if choose_ms else None,
                       "mrays_per_s": rays_total / (kernel_ms * 1e-3) / 1e6, "rays_per_path": rays_launch / paths_launch},
            "clocks": clocks,
            "e2e": {"value": paths_step * e2e_steps / (ms_e2e * 1e-3) / 1e6, "unit": "Mpaths/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / e2e_steps, "parts_ms": parts,
                    "host_frame": "pinned (spk_cuda_host_alloc)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic,
                         "dram_frac": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "note": "achieved = ALGORITHMIC bytes / time (SURVEY 8d). The working set of this scene is cache "
                                 "resident: dram_frac is the HBM utilisation, and ncu names the limiter (profiles/): issue "
                                 "slots and the CTA barrier of the sort, not memory" if key == "c2" else
                                 "achieved = ALGORITHMIC bytes / time (SURVEY 8d)",
                         "peak_source": peak_src,
                         "kernel": ("wf_extend_kernel<CULL=true> (+ wf_shade/regen/advance; achieved is over the whole step)"
                                    if use_wavefront else KERNEL_NAMES.get(form, f"megakernel form {form}")),
                         "alg_bytes_per_ray": alg_bytes / max(1.0, rays_launch),
                         "node_visits_per_ray": cst["node_visits"] / max(1, cst["rays"]),
                         "tri_tests_per_ray": cst["tri_tests"] / max(1, cst["rays"])},
        }
        if verify:
            out["multi_gpu_check"] = verify
        if cpu_baseline and world == 1:
            cpu_spp = 128 if scene == "testscene" else 4  # ~10 s on 15 workers (testscene)
            v, dt, cores, kind = cpu_render_rate(desc, cam, width, height, bounce, cpu_spp)
            out["cpu_baseline"] = {"value": v, "unit": "Mpaths/s", "cores": cores,
                                   "kind": "reference" if kind == "reference" else "port",
                                   "sample": f"{cpu_spp} spp of {spp} at {width}x{height} ({dt:.1f} s), linear in spp"}
    if group is not None:
        env.barrier()
        group.close()
    ctx.close()
    return out


def camera_grid_rays(torch, cam, n_side, dev):
    """n_side^2 COHERENT rays: the pinhole rays of an n_side x n_side pixel grid of `cam`, row-major (neighbours in the
    batch are neighbours on screen)"""
    def v(a):
        return torch.tensor(list(a), dtype=torch.float32, device=dev)
    pos, ll, mu, mv = v(cam.position), v(cam.lower_left), v(cam.max_u), v(cam.max_v)
    t = (torch.arange(n_side, dtype=torch.float32, device=dev) + 0.5) / n_side
    rays = torch.zeros((n_side * n_side, 8), dtype=torch.float32, device=dev)
    rays[:, 0:3] = pos
    for lo in range(0, n_side, 512):  # in slabs: the full (n, n, 3) temporary would be 800 MB
        hi = min(n_side, lo + 512)
        d = ll + t[None, :, None] * mu + t[lo:hi, None, None] * mv - pos
        d = d / d.norm(dim=2, keepdim=True)
        rays[lo * n_side:hi * n_side, 4:7] = d.reshape(-1, 3)
    return rays


def measure_trace(env, which, args, steps=None):
    """c3 (BASELINE.json configs[2]): soup-1M, 64 Mi incoherent rays, closest hit only; `terrain`: its coherent counterpart
    (terrain-1M, the camera rays of an 8192 x 8192 grid). One step = the whole batch; N ranks take contiguous slices, no
    collective. value: device-resident rays and hits; e2e: pinned host buffers through spk_cuda_trace_closest on a 16 Mi-ray
    slice of the batch (stated in config)."""
    from sparkle_b200 import _abi as A
    from sparkle_b200 import multi_gpu, scenes
    from sparkle_b200.device import CudaContext
    from sparkle_b200.scene import orbit_camera

    torch, rank, world, local = env.torch, env.rank, env.world, env.local
    steps = steps or args.steps
    warmup = max(3, args.warmup)
    total = 64 << 20
    begin, end = multi_gpu.ray_slice(total, rank, world)
    n = end - begin
    if which == "c3":
        sb, _ = scenes.soup(1_000_000)
        label = "soup-1M (1,000,000 triangles, one mesh), 64 Mi incoherent rays, closest hit only (BASELINE.json configs[2])"
    else:
        sb, camkw = scenes.terrain(1000, 500)
        label = ("terrain-1M (1,000,000 triangles, one mesh), 64 Mi COHERENT rays (camera rays of an 8192x8192 grid), "
                 "closest hit only (SURVEY 8d: C3's coherent counterpart)")
    desc = sb.finalize()
    ctx = CudaContext(64, 64, device=local)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.upload_scene(desc)
    if which == "c3":
        g = torch.Generator(device="cuda").manual_seed(2 + rank)
        rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
        rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 3.0 - 1.5
        d = torch.randn((n, 3), generator=g, device="cuda")
        rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
        del d
    else:
        rays = camera_grid_rays(torch, orbit_camera(8192, 8192, **camkw), 8192, env.dev)[begin:end].contiguous()
    hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")

    ctx.reset_stats()
    cn = min(n, 4 << 20)
    ctx.trace_closest_device(rays.data_ptr(), cn, hits.data_ptr(), A.FLAG_COUNTERS)
    cst = ctx.stats()
    sampler = ClockSampler(local, enabled=rank == 0).start()
    ms_total = env.timed(lambda: ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), 0), steps, warmup,
                         flush=False)
    clocks = sampler.stop()
    hit_frac = float((hits.view(torch.int32)[:, 3] != -1).float().mean().item())
    en = min(n, (16 << 20) // world)
    h_rays = rays[:en].cpu().pin_memory()  # pinned host buffers, as the contract asks
    h_hits = torch.empty((en, 5), dtype=torch.float32).pin_memory()

    def e2e_step():
        A.check(ctx.lib.spk_cuda_trace_closest(ctx.handle, h_rays.data_ptr(), en, h_hits.data_ptr(), 0))

    e2e_steps = max(1, min(steps, 3))
    ms_e2e = env.timed(e2e_step, e2e_steps, 1, flush=False)
    out = None
    if rank == 0:
        peak, peak_src = peaks()
        nv, tt = cst["node_visits"] / cn, cst["tri_tests"] / cn
        bytes_ray = 32 + 20 + 64 * nv + 48 * tt + 128 * cst["inst_visits"] / cn
        kernel_ms = ms_total / steps
        achieved = bytes_ray * n / (kernel_ms * 1e-3) / 1e9
        traffic = recorded_traffic(which, rays_per_launch=n) if world == 1 else None
        out = {"metric": "Mrays/s", "value": total * steps / (ms_total * 1e-3) / 1e6, "unit": "Mrays/s",
               "ms_per_step": kernel_ms, "steps": steps, "warmup": warmup,
               "config": {"workload": label, "parallelism": f"ray-slice x{world}",
                          "l2": "inputs larger than L2 (2.1 GB rays, 1.3 GB hits per step)",
                          "e2e_sample": f"{en * world} rays through pinned host buffers", "hit_fraction": hit_frac},
               "clocks": clocks,
               "e2e": {"value": en * world * e2e_steps / (ms_e2e * 1e-3) / 1e6, "unit": "Mrays/s",
                       "h2d_bytes_per_step": en * 32, "d2h_bytes_per_step": en * 20, "ms_per_step": ms_e2e / e2e_steps},
               "gpu_launches": steps,
               "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                            "traffic": traffic,
                            "dram_frac": (traffic / (kernel_ms * 1e-3) / 1e9 / peak) if traffic else None,
                            "peak_source": peak_src,
                            "kernel": "trace_persistent_kernel<CULL=true, VOTE=6>", "alg_bytes_per_ray": bytes_ray,
                            "node_visits_per_ray": nv, "tri_tests_per_ray": tt}}
    del rays, hits
    ctx.close()
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + ["c3", "terrain"])
    ap.add_argument("--spp", type=int, default=0, help="bound the workload's spp (stated in config.workload)")
    ap.add_argument("--tracer", default="auto", choices=["auto", "megakernel", "wavefront"])
    ap.add_argument("--reduce", default="group", choices=["group", "nccl"],
                    help="N > 1: the library's fused peer-memory reduce + resolve (default) or one ncclReduce (A/B)")
    ap.add_argument("--extra", default="c3,terrain,c4,c5",
                    help="other BASELINE.json configs measured into the same JSON line ('' or 'none' = headline only)")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.workload in ("c3", "terrain"):
            raise SystemExit("bench.py: the ray-batch workloads have no reference arm here (tests compare them with the oracle)")
        return run_reference(args)
    env = Env()
    head = {"higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "n_gpus": env.world}
    if args.workload in ("c3", "terrain"):
        res = measure_trace(env, args.workload, args)
    else:
        res = measure_render(env, args.workload, args, spp_bound=args.spp, cpu_baseline=not args.no_cpu_baseline)
    extra = {}
    for key in [k for k in args.extra.split(",") if k and k != "none" and k != args.workload]:
        esteps = max(2, min(args.steps, 3))
        if key in ("c3", "terrain"):
            r = measure_trace(env, key, args, steps=esteps)
        else:
            r = measure_render(env, key, args, spp_bound=WORKLOADS[key][5], steps=esteps)
        if r is not None:
            extra[key] = r
    if env.rank == 0:
        line = {"metric": res.pop("metric", "Mpaths/s")}
        line.update(head)
        line.update(res)
        if extra:
            line["extra"] = extra
        order = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step"]
        print(json.dumps({**{k: line[k] for k in order if k in line}, **{k: v for k, v in line.items() if k not in order}}),
              flush=True)
    env.close()


if __name__ == "__main__":
    main()
