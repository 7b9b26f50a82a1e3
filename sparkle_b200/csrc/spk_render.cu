/*
 * spk_render.cu — the path tracer (persistent-thread megakernel form) and the resolve (tone-mapping) pass.
 *
 *   CPURenderer::BasePass     libraries/source/renderer/renderer/CPURenderer.cpp:419-450  per-pixel reseed, one task per row
 *   CPURenderer::RenderPixel  :398-417   jitter (2 draws), SamplePixel, clamp to OutputLimit, (rgb, valid) out
 *   SetupViewRay              :220-238   thin lens, UnitDisk always drawn
 *   SamplePixel               :240-396   bounce loop, sky on miss, emissive stop, early-out, Russian roulette from bounce 3
 *   DenoisePass (temporal)    :530-544   acc = lerp(new, acc, N/(N+spp))
 *   ToneMappingPass/ACESFilm  :554-576
 *
 * Work decomposition: one thread owns one pixel and loops over the samples of this call, keeping the pixel's
 * accumulator in registers (one 16-byte read-modify-write of the accumulation buffer per pixel per call, not per
 * sample). A warp covers an 8x4 pixel tile so primary rays stay coherent. Sample s of pixel p depends only on (p, s)
 * through its seed, so any split of the sample range over launches or GPUs gives the same sum.
 * Compiled with -fmad=false (spk_device.cuh).
 */
#include "spk_context.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>
#include "spk_shade.cuh"
#include "spk_traverse.cuh"

namespace spk
{
namespace
{
constexpr int kTileW = 8, kTileH = 4; // one warp
constexpr int kWarpsPerBlock = 4;

struct RenderArgs
{
    uint32_t width, height;
    uint32_t first_sample, sample_count, sample_stride, max_bounce;
    float output_limit;
    uint32_t accum_mode;
    uint32_t debug;
    uint32_t cumulated; // samples already in the accumulator (running-average mode)
    uint32_t cumulated_cap; // the camera's cap on that count (max_spp, CameraRenderProxy.cpp:70-73); 0xFFFFFFFF none
    uint32_t fresh;         // 1: the accumulator was reset and not cleared yet - start from zero instead of reading it
};

// GBUF: one sample per pixel into a g-buffer instead of the accumulator (RenderPixel with spatial_denoise,
// CPURenderer.cpp:411-416): gbuf[p] = (min(colour, limit), valid), gbuf[W*H + p] = (first-hit normal, 0)
template <bool CULL, bool COUNT, bool GBUF = false>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
    render_kernel(const DScene sc, const DCamera cam, const RenderArgs a, float4 *__restrict__ accum,
                  unsigned long long *__restrict__ counters)
{
    const uint32_t tiles_x = (a.width + kTileW - 1) / kTileW;
    const uint32_t warp = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t i = (warp % tiles_x) * kTileW + (lane & (kTileW - 1));
    const uint32_t j = (warp / tiles_x) * kTileH + (lane / kTileW);
    uint32_t rays = 0;
    TraceCounters cnt{0, 0, 0, 0};
    if (i < a.width && j < a.height)
    {
        const float pw = 1.f / (float)(a.width - 1), ph = 1.f / (float)(a.height - 1);
        const V3 cpos = mk(cam.position[0], cam.position[1], cam.position[2]);
        const V3 cup = mk(cam.up[0], cam.up[1], cam.up[2]);
        const V3 cright = mk(cam.right[0], cam.right[1], cam.right[2]);
        const V3 max_u = mk(cam.max_u[0], cam.max_u[1], cam.max_u[2]);
        const V3 max_v = mk(cam.max_v[0], cam.max_v[1], cam.max_v[2]);
        const V3 lower_left = mk(cam.lower_left[0], cam.lower_left[1], cam.lower_left[2]);
        float4 acc = (GBUF || a.fresh) ? make_float4(0.f, 0.f, 0.f, 0.f) : accum[size_t(j) * a.width + i];
        V3 first_normal = mk(0.f, 0.f, 0.f);
        uint32_t stack[kThreadStack];
        for (uint32_t s = 0; s < a.sample_count; s++)
        {
            // CPURenderer.cpp:442-443, unsigned wrap
            const uint32_t frame_seed = a.first_sample + s * a.sample_stride;
            Rng rng = rng_seed(j * a.width + i + frame_seed * a.width * a.height);
            float u = ((float)i + rng_unit(rng)) * pw;
            float v = ((float)j + rng_unit(rng)) * ph;
            // SetupViewRay
            float dx, dy;
            unit_disk(rng, dx, dy);
            dx = dx * cam.aperture_radius;
            dy = dy * cam.aperture_radius;
            V3 origin = cpos + (dx * cright + dy * cup);
            V3 target = lower_left + u * max_u + v * max_v;
            V3 dir = normalized(target - origin);

            V3 throughput = mk(1.f, 1.f, 1.f), color = mk(0.f, 0.f, 0.f);
            float valid = 1.f;
            for (uint32_t bounce = 0; bounce < a.max_bounce; bounce++)
            {
                Hit hit;
                trace_closest<CULL, COUNT>(sc, origin, dir, hit, stack, cnt);
                rays++;
                if (hit.inst == SPK_INVALID_ID)
                {
                    if (sc.sky.has_sky)
                    {
                        color = color + cmul(sky_evaluate(sc, dir), throughput);
                    }
                    break;
                }
                Surface sf;
                get_intersection(sc, origin, dir, hit, sf);
                if (GBUF && bounce == 0)
                {
                    first_normal = sf.normal; // SampleResult::world_normal (CPURenderer.cpp:280-283)
                }
                if (a.debug)
                {
                    // parity probes (tests only): first-hit normal / tangent / (u, v, t)
                    color = a.debug == 1 ? sf.normal : a.debug == 2 ? sf.tangent : mk(sf.u, sf.v, hit.t);
                    break;
                }
                V3 emissive = material_emissive(sc, sc.materials[sf.material], sf.u, sf.v);
                if (norm(emissive) > kEps)
                {
                    color = color + cmul(emissive, throughput);
                    break;
                }
                V3 next_dir;
                V3 this_throughput = sample_surface(sc, rng, sf, dir, next_dir);
                throughput = cmul(throughput, this_throughput);
                if (sqnorm(throughput) < kEps)
                {
                    valid = -1.f;
                    break;
                }
                if (bounce >= 3)
                {
                    float p = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
                    p = p < 0.05f ? 0.05f : (p > 1.0f ? 1.0f : p); // std::clamp
                    if (rng_unit(rng) > p)
                    {
                        break;
                    }
                    throughput = throughput / p;
                }
                origin = sf.location + next_dir * kTolerance;
                dir = next_dir;
            }
            color = mk(fminf(color.x, a.output_limit), fminf(color.y, a.output_limit), fminf(color.z, a.output_limit));
            if (GBUF)
            {
                acc = make_float4(color.x, color.y, color.z, valid);
            }
            else if (a.accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
            {
                uint32_t n = min(a.cumulated + s, a.cumulated_cap);
                float ma = (float)n / (float)(n + 1u);
                acc = make_float4(lerp1(color.x, acc.x, ma), lerp1(color.y, acc.y, ma), lerp1(color.z, acc.z, ma),
                                  lerp1(valid, acc.w, ma));
            }
            else
            {
                acc = make_float4(acc.x + color.x, acc.y + color.y, acc.z + color.z, acc.w + valid);
            }
        }
        accum[size_t(j) * a.width + i] = acc;
        if (GBUF)
        {
            accum[size_t(a.width) * a.height + size_t(j) * a.width + i] =
                make_float4(first_normal.x, first_normal.y, first_normal.z, 0.f);
        }
    }
    uint32_t total = __reduce_add_sync(0xffffffffu, rays);
    uint32_t pixels = __popc(__ballot_sync(0xffffffffu, i < a.width && j < a.height));
    if (lane == 0 && pixels)
    {
        atomicAdd(counters + 0, (unsigned long long)pixels * a.sample_count);
        atomicAdd(counters + 1, (unsigned long long)total);
    }
    if (COUNT)
    {
        uint32_t v[4] = {cnt.node_visits, cnt.tri_tests, cnt.inst_visits, cnt.sphere_tests};
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            uint32_t sum = __reduce_add_sync(0xffffffffu, v[k]);
            if (lane == 0 && sum)
            {
                atomicAdd(counters + 2 + k, (unsigned long long)sum);
            }
        }
    }
}


// ---- phased form: the same per-pixel loop, but the CTA walks through it in lock step ----
// ncu on render_kernel: 64-68 % of warp-stall samples are `no_instruction` — 28 warps per SM, each somewhere else in a
// 66 KiB program, against a 32 KiB L1.5 instruction cache. Here every iteration has three phases separated by CTA
// barriers — (A) start the next sample where a path just ended, (B) trace one segment, (C) shade it — so all warps of
// a CTA fetch the same few KiB at the same time. Per-pixel arithmetic and the order samples are added are unchanged.
template <bool CULL, bool COUNT, int WARPS, int MINB, bool SYNC2>
__global__ void __launch_bounds__(WARPS * 32, MINB)
    render_phased_kernel(const DScene sc, const DCamera cam, const RenderArgs a, float4 *__restrict__ accum,
                         unsigned long long *__restrict__ counters)
{
    const uint32_t tiles_x = (a.width + kTileW - 1) / kTileW;
    const uint32_t warp = blockIdx.x * WARPS + (threadIdx.x >> 5);
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t i = (warp % tiles_x) * kTileW + (lane & (kTileW - 1));
    const uint32_t j = (warp / tiles_x) * kTileH + (lane / kTileW);
    const bool in_image = i < a.width && j < a.height;
    uint32_t rays = 0;
    TraceCounters cnt{0, 0, 0, 0};
    const float pw = 1.f / (float)(a.width - 1), ph = 1.f / (float)(a.height - 1);
    const V3 cpos = mk(cam.position[0], cam.position[1], cam.position[2]);
    const V3 cup = mk(cam.up[0], cam.up[1], cam.up[2]);
    const V3 cright = mk(cam.right[0], cam.right[1], cam.right[2]);
    const V3 max_u = mk(cam.max_u[0], cam.max_u[1], cam.max_u[2]);
    const V3 max_v = mk(cam.max_v[0], cam.max_v[1], cam.max_v[2]);
    const V3 lower_left = mk(cam.lower_left[0], cam.lower_left[1], cam.lower_left[2]);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_image)
    {
        acc = a.fresh ? make_float4(0.f, 0.f, 0.f, 0.f) : accum[size_t(j) * a.width + i];
    }
    uint32_t stack[kThreadStack];
    uint32_t s = 0, bounce = 0;
    bool live = in_image, need_path = true;
    Rng rng;
    V3 origin, dir, throughput, color;
    float valid = 1.f;
    for (;;)
    {
        // ---- phase A: next sample of this pixel
        if (live && need_path)
        {
            if (s >= a.sample_count)
            {
                live = false;
            }
            else
            {
                // CPURenderer.cpp:442-443, unsigned wrap
                const uint32_t frame_seed = a.first_sample + s * a.sample_stride;
                rng = rng_seed(j * a.width + i + frame_seed * a.width * a.height);
                float u = ((float)i + rng_unit(rng)) * pw;
                float v = ((float)j + rng_unit(rng)) * ph;
                // SetupViewRay
                float dx, dy;
                unit_disk(rng, dx, dy);
                dx = dx * cam.aperture_radius;
                dy = dy * cam.aperture_radius;
                origin = cpos + (dx * cright + dy * cup);
                V3 target = lower_left + u * max_u + v * max_v;
                dir = normalized(target - origin);
                throughput = mk(1.f, 1.f, 1.f);
                color = mk(0.f, 0.f, 0.f);
                valid = 1.f;
                bounce = 0;
                need_path = false;
            }
        }
        if (!__syncthreads_or(live))
        {
            break;
        }
        // ---- phase B: one path segment
        bool done = false;
        Hit hit;
        hit.inst = SPK_INVALID_ID;
        if (live && a.max_bounce != 0u)
        {
            trace_closest<CULL, COUNT>(sc, origin, dir, hit, stack, cnt);
            rays++;
        }
        if (SYNC2)
        {
            __syncthreads();
        }
        // ---- phase C: shade it
        if (live)
        {
            if (a.max_bounce == 0u)
            {
                done = true; // a zero-iteration bounce loop: the sample is black and valid
            }
            else if (hit.inst == SPK_INVALID_ID)
            {
                if (sc.sky.has_sky)
                {
                    color = color + cmul(sky_evaluate(sc, dir), throughput);
                }
                done = true;
            }
            else
            {
                Surface sf;
                get_intersection(sc, origin, dir, hit, sf);
                V3 emissive = material_emissive(sc, sc.materials[sf.material], sf.u, sf.v);
                if (a.debug)
                {
                    // parity probes (tests only): first-hit normal / tangent / (u, v, t)
                    color = a.debug == 1 ? sf.normal : a.debug == 2 ? sf.tangent : mk(sf.u, sf.v, hit.t);
                    done = true;
                }
                else if (norm(emissive) > kEps)
                {
                    color = color + cmul(emissive, throughput);
                    done = true;
                }
                else
                {
                    V3 next_dir;
                    V3 this_throughput = sample_surface(sc, rng, sf, dir, next_dir);
                    throughput = cmul(throughput, this_throughput);
                    if (sqnorm(throughput) < kEps)
                    {
                        valid = -1.f;
                        done = true;
                    }
                    else
                    {
                        if (bounce >= 3)
                        {
                            float p = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
                            p = p < 0.05f ? 0.05f : (p > 1.0f ? 1.0f : p); // std::clamp
                            if (rng_unit(rng) > p)
                            {
                                done = true;
                            }
                            else
                            {
                                throughput = throughput / p;
                            }
                        }
                        if (!done)
                        {
                            origin = sf.location + next_dir * kTolerance;
                            dir = next_dir;
                            bounce++;
                            done = bounce >= a.max_bounce; // the bounce cap adds nothing (CPURenderer.cpp:256)
                        }
                    }
                }
            }
            if (done)
            {
                color = mk(fminf(color.x, a.output_limit), fminf(color.y, a.output_limit), fminf(color.z, a.output_limit));
                if (a.accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
                {
                    uint32_t n = min(a.cumulated + s, a.cumulated_cap);
                    float ma = (float)n / (float)(n + 1u);
                    acc = make_float4(lerp1(color.x, acc.x, ma), lerp1(color.y, acc.y, ma), lerp1(color.z, acc.z, ma),
                                      lerp1(valid, acc.w, ma));
                }
                else
                {
                    acc = make_float4(acc.x + color.x, acc.y + color.y, acc.z + color.z, acc.w + valid);
                }
                s++;
                need_path = true;
            }
        }
    }
    if (in_image)
    {
        accum[size_t(j) * a.width + i] = acc;
    }
    uint32_t total = __reduce_add_sync(0xffffffffu, rays);
    uint32_t pixels = __popc(__ballot_sync(0xffffffffu, in_image));
    if (lane == 0 && pixels)
    {
        atomicAdd(counters + 0, (unsigned long long)pixels * a.sample_count);
        atomicAdd(counters + 1, (unsigned long long)total);
    }
    if (COUNT)
    {
        uint32_t v[4] = {cnt.node_visits, cnt.tri_tests, cnt.inst_visits, cnt.sphere_tests};
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            uint32_t sum = __reduce_add_sync(0xffffffffu, v[k]);
            if (lane == 0 && sum)
            {
                atomicAdd(counters + 2 + k, (unsigned long long)sum);
            }
        }
    }
}

// ---- sorted form: the phased kernel, with the CTA's segments regrouped by material class before shading ----
// ncu on render_phased_kernel: shading runs with 11.2 of 32 lanes (profiles/r1_render_phased_ncu.txt) because the
// 256 pixels of a CTA hit different materials. Here phase C goes through shared memory: every thread parks its segment
// (ray, throughput, rng, hit: 19 words) in a [word][thread] array, the CTA counting-sorts the 256 segments by
// shade_class (dead lanes last, so they idle as whole warps), thread k shades the k-th segment of the sorted order and
// writes the outcome back into the owner's column, and the owner picks it up and accumulates exactly as before: the
// per-pixel arithmetic and the order samples are added are unchanged.
// SPK_SMEM_SLAB 1: a worker's world-space slab constants in six more shared-memory words per thread instead of seven local
// stores at every instance entry. Measured (B200, 1080p, 8x4): room 32 spp 117.4 -> 115.7 ms, TestScene 64 spp 31.44 ->
// 31.80 ms; off, the headline scene decides.
#ifndef SPK_SMEM_SLAB
#define SPK_SMEM_SLAB 0
#endif
template <int WARPS> constexpr size_t sorted_smem_bytes()
{
    return size_t((23 + (SPK_SMEM_SLAB ? 6 : 0)) * WARPS * 32) * 4; // path slots (+ a worker's world-space slab constants)
}

// Measured and dropped: a CTA-local ray QUEUE for phase B (the 256 parked rays drained by the warps with
// warp_trace_loop, lanes refilling from a shared-memory cursor). With only 256 rays per phase the loop's start-up, votes
// and drain cost more than the idle lanes they save: TestScene 36.5 vs 23.4 ms per 32 spp, city 415 vs 327, room 233 vs
// 159 (B200, same launch shapes).
// PULL: the grid is persistent (SM count x CTAs per SM) and a column is not tied to a pixel. When a column's pixel has
// all its samples the thread that shaded its last segment writes the pixel's sum to the accumulation buffer and takes the
// NEXT PIXEL from the CTA's pixel pool, so columns stay busy until the image is exhausted. Without it a CTA owns 512
// pixels and lives until its SLOWEST pixel has finished its samples (a pixel on glass needs up to 8x the segments of a
// sky pixel), holding its half of the SM's threads and shared memory with a few live columns.
// The pool is two chunks of T tile-linear pixel indices (32 consecutive indices are one 8x4 tile, T of them one strip
// of tiles, the footprint of an owner-form CTA, so the CTA's rays stay as coherent as there) in shared memory; thread 0
// swaps a fresh chunk in between the barriers of the sort when the first runs dry and keeps the NEXT chunk's global
// atomicAdd in flight across a whole iteration. The first pulling form took every pixel with its own global atomicAdd
// from inside the shading phase and then read the old accumulator value: two dependent ~1 us round trips on the CTA's
// critical path in almost every iteration, and pixels from all over the image in one CTA - it measured 15 % SLOWER
// than the owner form (TestScene 21.7 vs 18.8 ms per 32 spp). A pixel's samples are still added in order by one thread
// at a time, so the image is bit-identical to the owner forms. The column's pixel index lives in 16 bits of shared
// memory plus 7 bits of the packed word (sample : 17 | pixel_hi : 7 | bounce : 8): < 2^23 pixels, < 2^17 samples per call.
// STABLE (MODE 2): the counting sort keeps the columns of a class in pixel order. Thread t looks at column t's class after
// the trace: the lanes of a warp that agree (__match_any_sync) are one 8x4 tile's segments of that class, they take a
// contiguous run of the class with ONE shared-memory atomicAdd by their first lane, and keep their order inside it. A warp
// of the sorted order is then a few runs of tile neighbours: the rays it traces next are coherent (camera rays of
// adjacent pixels, bounces off the same surface), its shared-memory columns are consecutive (the arbitrary order of the
// per-thread atomicAdd form cost 2 of every 3 shared-memory wavefronts in bank conflicts), at the price of one more
// barrier per iteration.
template <bool CULL, bool COUNT, int WARPS, int MINB, int MODE>
__global__ void __launch_bounds__(WARPS * 32, MINB)
    render_sorted_kernel(const DScene sc, const DCamera cam, const RenderArgs a, float4 *__restrict__ accum,
                         unsigned long long *__restrict__ counters, const uint32_t *__restrict__ cta_order,
                         uint32_t *__restrict__ cta_cost)
{
    constexpr bool PULL = MODE == 1, STABLE = MODE == 2;
    // Owner forms: which strip of tiles this CTA renders. The hardware hands CTAs out in blockIdx order; cta_order (when
    // given) lists the strips by DECREASING cost of the previous launch of the same view, so the long strips (glass,
    // deep meshes) start first and the last wave is made of short ones (longest-processing-time-first: the tail of the
    // grid was ~6 % of the launch). cta_cost receives this launch's cycles per strip for the next one.
    const long long t_start = clock64();
    const uint32_t strip = (!PULL && cta_order) ? cta_order[blockIdx.x] : blockIdx.x;
    constexpr int T = WARPS * 32;
    __shared__ uint8_t s_key[STABLE ? T : 1]; // STABLE: the shading class of every column's parked segment, 0xFF = no path
    extern __shared__ __align__(16) unsigned char s_dynamic[]; // sorted_smem_bytes<WARPS>()
    float *s_state = reinterpret_cast<float *>(s_dynamic);
    __shared__ uint32_t s_bin[2][32];
    __shared__ uint32_t s_total;
    __shared__ uint16_t s_sorted[T];
    __shared__ uint16_t s_pixel_lo[PULL ? T : 1];
    const uint32_t tid = threadIdx.x;
    const uint32_t tiles_x = (a.width + kTileW - 1) / kTileW;
    const uint32_t lane = tid & 31;
    constexpr uint32_t kSampleShift = PULL ? 15u : 8u; // packed word: sample above, (pixel_hi,) bounce below
    // column c of the CTA <-> pixel: a warp-sized run of columns is one 8x4 tile
    auto pixel_of = [&](uint32_t c, uint32_t &pi, uint32_t &pj) {
        const uint32_t w = strip * WARPS + (c >> 5), l = c & 31u;
        pi = (w % tiles_x) * kTileW + (l & (kTileW - 1));
        pj = (w / tiles_x) * kTileH + (l / kTileW);
    };
    // PULL: tile-linear pixel index <-> pixel
    auto pixel_of_index = [&](uint32_t p, uint32_t &pi, uint32_t &pj) {
        const uint32_t w = p >> 5, l = p & 31u;
        pi = (w % tiles_x) * kTileW + (l & (kTileW - 1));
        pj = (w / tiles_x) * kTileH + (l / kTileW);
    };
    const uint32_t total_px = tiles_x * ((a.height + kTileH - 1) / kTileH) * 32u;
    unsigned int *next_pixel = reinterpret_cast<unsigned int *>(counters + 7);
    uint32_t i = 0, j = 0;
    bool in_image;
    uint32_t pixels_done = 0;
    __shared__ uint32_t s_pool[PULL ? 4 : 1]; // PULL: {next, end} of the chunk in use and of the reserve chunk
    uint32_t prefetched = 0;                  // PULL, thread 0: base of the chunk after those, its atomicAdd in flight
    auto clamp_px = [&](uint32_t p) { return p < total_px ? p : total_px; };
    if (PULL)
    {
        // three chunks to start with: one for the columns themselves, two for the pool
        if (tid == 0)
        {
            const uint32_t base = atomicAdd(next_pixel, 3u * T);
            s_pool[0] = clamp_px(base + T);
            s_pool[1] = clamp_px(base + 2u * T);
            s_pool[2] = clamp_px(base + 2u * T);
            s_pool[3] = clamp_px(base + 3u * T);
            s_total = base;
            prefetched = atomicAdd(next_pixel, (uint32_t)T);
        }
        __syncthreads();
        const uint32_t p = s_total + tid;
        in_image = false;
        if (p < total_px)
        {
            pixel_of_index(p, i, j);
            in_image = i < a.width && j < a.height;
        }
        if (in_image)
        {
            s_pixel_lo[tid] = (uint16_t)p;
            s_state[tid + 18 * T] = __uint_as_float((p >> 16) << 8); // pixel_hi; start_path keeps it
        }
    }
    else
    {
        pixel_of(tid, i, j);
        in_image = i < a.width && j < a.height;
    }
    uint32_t rays = 0;
    TraceCounters cnt{0, 0, 0, 0};
    const float pw = 1.f / (float)(a.width - 1), ph = 1.f / (float)(a.height - 1);
    uint32_t stack[kThreadStack];
    // Every column of the shared-memory array is one pixel's path slot: ray 0-5, throughput 6-8, rng 9-12, hit 13-17,
    // (sample index << 8 | bounce) 18, the pixel's accumulator 19-22. Threads are WORKERS, not owners: after the sort,
    // thread k shades the k-th segment, adds a finished sample to the column's accumulator, starts the pixel's next
    // sample there, and goes straight on to trace the ray it has just produced - no barrier between shading and
    // tracing, so a CTA waits once per segment (before the sort), not twice. One column is touched by one thread per
    // iteration and a pixel's samples end in order, so the sums are those of the owner-thread forms, bit for bit.
    auto start_path = [&](float *c, uint32_t pi, uint32_t pj, uint32_t sample) {
        const uint32_t frame_seed = a.first_sample + sample * a.sample_stride; // CPURenderer.cpp:442-443
        Rng rng = rng_seed(pj * a.width + pi + frame_seed * a.width * a.height);
        float u = ((float)pi + rng_unit(rng)) * pw;
        float v = ((float)pj + rng_unit(rng)) * ph;
        float dx, dy;
        unit_disk(rng, dx, dy);
        dx = dx * cam.aperture_radius;
        dy = dy * cam.aperture_radius;
        const V3 cpos = mk(cam.position[0], cam.position[1], cam.position[2]);
        const V3 cup = mk(cam.up[0], cam.up[1], cam.up[2]);
        const V3 cright = mk(cam.right[0], cam.right[1], cam.right[2]);
        const V3 max_u = mk(cam.max_u[0], cam.max_u[1], cam.max_u[2]);
        const V3 max_v = mk(cam.max_v[0], cam.max_v[1], cam.max_v[2]);
        const V3 lower_left = mk(cam.lower_left[0], cam.lower_left[1], cam.lower_left[2]);
        const V3 origin = cpos + (dx * cright + dy * cup);
        V3 target = lower_left + u * max_u + v * max_v;
        const V3 dir = normalized(target - origin);
        c[0 * T] = origin.x;
        c[1 * T] = origin.y;
        c[2 * T] = origin.z;
        c[3 * T] = dir.x;
        c[4 * T] = dir.y;
        c[5 * T] = dir.z;
        c[6 * T] = 1.f; // throughput
        c[7 * T] = 1.f;
        c[8 * T] = 1.f;
        c[9 * T] = __uint_as_float(rng.s0);
        c[10 * T] = __uint_as_float(rng.s1);
        c[11 * T] = __uint_as_float(rng.s2);
        c[12 * T] = __uint_as_float(rng.s3);
        // sample index | (PULL: the column's pixel_hi, kept) | bounce = 0
        c[18 * T] = __uint_as_float((sample << kSampleShift) | (PULL ? (__float_as_uint(c[18 * T]) & 0x7F00u) : 0u));
    };
    float *mine = s_state + tid; // the column this thread works on in the current iteration
    bool have = in_image && a.sample_count > 0u;
    if (STABLE)
    {
        s_key[tid] = 0xFFu;
    }
    if (in_image)
    {
        const float4 acc = a.fresh ? make_float4(0.f, 0.f, 0.f, 0.f) : accum[size_t(j) * a.width + i];
        mine[19 * T] = acc.x;
        mine[20 * T] = acc.y;
        mine[21 * T] = acc.z;
        mine[22 * T] = acc.w;
    }
    if (have)
    {
        start_path(mine, i, j, 0u);
    }
    if (tid < 64)
    {
        s_bin[tid >> 5][tid & 31u] = 0;
    }
    __syncthreads();
    uint32_t parity = 0;
    for (;;)
    {
        // ---- trace the segment in this thread's column, park the hit, count its shading class
        uint32_t key = 0, pos = 0;
        if (have)
        {
            Hit hit;
            const V3 origin = mk(mine[0 * T], mine[1 * T], mine[2 * T]), dir = mk(mine[3 * T], mine[4 * T], mine[5 * T]);
            trace_closest<CULL, COUNT, SPK_SMEM_SLAB != 0>(sc, origin, dir, hit, stack, cnt, s_state + 23 * T + tid, T);
            rays++;
            mine[13 * T] = hit.t;
            mine[14 * T] = hit.u;
            mine[15 * T] = hit.v;
            mine[16 * T] = __uint_as_float(hit.inst);
            mine[17 * T] = __uint_as_float(hit.face);
            key = shade_class(sc, hit.inst);
            if (STABLE)
            {
                s_key[mine - s_state] = (uint8_t)key;
            }
            else
            {
                pos = atomicAdd(&s_bin[parity][key], 1u);
            }
        }
        __syncthreads();
        uint32_t sort_column = uint32_t(mine - s_state);
        if (STABLE)
        {
            // from here to the scatter thread t speaks for COLUMN t: tile neighbours of one class share an atomicAdd
            key = s_key[tid];
            have = key != 0xFFu;
            sort_column = tid;
            const uint32_t same = __match_any_sync(0xffffffffu, key);
            const uint32_t rank = __popc(same & ((1u << lane) - 1u));
            uint32_t base = 0;
            if (have && rank == 0u)
            {
                base = atomicAdd(&s_bin[parity][key], (uint32_t)__popc(same));
            }
            pos = __shfl_sync(0xffffffffu, base, __ffs(same) - 1) + rank;
            key &= 31u;
            __syncthreads();
        }
        // every warp scans the 32 class counts itself (5 shuffles) instead of waiting for warp 0 behind another barrier
        uint32_t n_live;
        {
            const uint32_t c = s_bin[parity][lane];
            uint32_t incl = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)lane >= d)
                {
                    incl += t;
                }
            }
            n_live = __shfl_sync(0xffffffffu, incl, 31);
            const uint32_t start = __shfl_sync(0xffffffffu, incl - c, key); // exclusive start of this thread's class
            if (have)
            {
                s_sorted[start + pos] = (uint16_t)sort_column;
            }
        }
        if (tid < 32)
        {
            s_bin[parity ^ 1u][tid] = 0; // the other set was last read before the previous iteration's last barrier
        }
        if (PULL && tid == 0 && s_pool[0] >= s_pool[1])
        {
            // nobody takes pixels between the two barriers: the reserve chunk moves up, the prefetched one becomes the
            // reserve, and the chunk after it is asked for now and used an iteration or more later
            s_pool[0] = s_pool[2] < s_pool[3] ? s_pool[2] : s_pool[3];
            s_pool[1] = s_pool[3];
            s_pool[2] = clamp_px(prefetched);
            s_pool[3] = clamp_px(prefetched + T);
            if (prefetched < total_px)
            {
                prefetched = atomicAdd(next_pixel, (uint32_t)T);
            }
        }
        if (n_live == 0u)
        {
            break;
        }
        __syncthreads();
        parity ^= 1u;
        have = tid < n_live;
        if (have)
        {
            const uint32_t column = s_sorted[tid];
            float *src = s_state + column;
            mine = src;
            V3 o = mk(src[0 * T], src[1 * T], src[2 * T]), d = mk(src[3 * T], src[4 * T], src[5 * T]);
            V3 thr = mk(src[6 * T], src[7 * T], src[8 * T]);
            Hit h;
            h.t = src[13 * T];
            h.u = src[14 * T];
            h.v = src[15 * T];
            h.inst = __float_as_uint(src[16 * T]);
            h.face = __float_as_uint(src[17 * T]);
            const uint32_t packed = __float_as_uint(src[18 * T]);
            const uint32_t sample = packed >> kSampleShift;
            uint32_t b = packed & 255u;
            V3 color = mk(0.f, 0.f, 0.f);
            float valid = 1.f;
            bool done = false;
            if (h.inst == SPK_INVALID_ID)
            {
                if (sc.sky.has_sky)
                {
                    color = cmul(sky_evaluate(sc, d), thr);
                }
                done = true;
            }
            else
            {
                Surface sf;
                get_intersection(sc, o, d, h, sf);
                V3 emissive = material_emissive(sc, sc.materials[sf.material], sf.u, sf.v);
                if (norm(emissive) > kEps)
                {
                    color = cmul(emissive, thr);
                    done = true;
                }
                else
                {
                    Rng r{__float_as_uint(src[9 * T]), __float_as_uint(src[10 * T]), __float_as_uint(src[11 * T]),
                          __float_as_uint(src[12 * T])};
                    V3 next_dir;
                    V3 this_throughput = sample_surface(sc, r, sf, d, next_dir);
                    thr = cmul(thr, this_throughput);
                    if (sqnorm(thr) < kEps)
                    {
                        valid = -1.f;
                        done = true;
                    }
                    else
                    {
                        if (b >= 3)
                        {
                            float p = fmaxf(thr.x, fmaxf(thr.y, thr.z));
                            p = p < 0.05f ? 0.05f : (p > 1.0f ? 1.0f : p); // std::clamp
                            if (rng_unit(r) > p)
                            {
                                done = true;
                            }
                            else
                            {
                                thr = thr / p;
                            }
                        }
                        if (!done)
                        {
                            b++;
                            done = b >= a.max_bounce; // the bounce cap adds nothing (CPURenderer.cpp:256)
                        }
                        if (!done)
                        {
                            V3 no = sf.location + next_dir * kTolerance;
                            src[0 * T] = no.x;
                            src[1 * T] = no.y;
                            src[2 * T] = no.z;
                            src[3 * T] = next_dir.x;
                            src[4 * T] = next_dir.y;
                            src[5 * T] = next_dir.z;
                            src[6 * T] = thr.x;
                            src[7 * T] = thr.y;
                            src[8 * T] = thr.z;
                            src[9 * T] = __uint_as_float(r.s0);
                            src[10 * T] = __uint_as_float(r.s1);
                            src[11 * T] = __uint_as_float(r.s2);
                            src[12 * T] = __uint_as_float(r.s3);
                            src[18 * T] = __uint_as_float((packed & ~255u) | b);
                        }
                    }
                }
            }
            if (done)
            {
                color = mk(fminf(color.x, a.output_limit), fminf(color.y, a.output_limit), fminf(color.z, a.output_limit));
                float4 acc = make_float4(src[19 * T], src[20 * T], src[21 * T], src[22 * T]);
                if (a.accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
                {
                    uint32_t n = min(a.cumulated + sample, a.cumulated_cap);
                    float ma = (float)n / (float)(n + 1u);
                    acc = make_float4(lerp1(color.x, acc.x, ma), lerp1(color.y, acc.y, ma), lerp1(color.z, acc.z, ma),
                                      lerp1(valid, acc.w, ma));
                }
                else
                {
                    acc = make_float4(acc.x + color.x, acc.y + color.y, acc.z + color.z, acc.w + valid);
                }
                src[19 * T] = acc.x;
                src[20 * T] = acc.y;
                src[21 * T] = acc.z;
                src[22 * T] = acc.w;
                if (sample + 1u < a.sample_count)
                {
                    uint32_t pi, pj;
                    if (PULL)
                    {
                        pixel_of_index((((packed >> 8) & 0x7Fu) << 16) | s_pixel_lo[column], pi, pj);
                    }
                    else
                    {
                        pixel_of(column, pi, pj);
                    }
                    start_path(src, pi, pj, sample + 1u);
                }
                else if (PULL)
                {
                    // the pixel has all its samples: store its sum, take the next pixel of the image for this column
                    uint32_t pi, pj;
                    pixel_of_index((((packed >> 8) & 0x7Fu) << 16) | s_pixel_lo[column], pi, pj);
                    accum[size_t(pj) * a.width + pi] = acc;
                    pixels_done++;
                    have = false;
                    for (;;)
                    {
                        uint32_t p = atomicAdd(&s_pool[0], 1u);
                        if (p >= s_pool[1])
                        {
                            p = atomicAdd(&s_pool[2], 1u);
                            if (p >= s_pool[3])
                            {
                                break; // image exhausted: the column stays out of the next sort
                            }
                        }
                        pixel_of_index(p, pi, pj);
                        if (pi < a.width && pj < a.height)
                        {
                            const float4 old = a.fresh ? make_float4(0.f, 0.f, 0.f, 0.f) : accum[size_t(pj) * a.width + pi];
                            src[19 * T] = old.x;
                            src[20 * T] = old.y;
                            src[21 * T] = old.z;
                            src[22 * T] = old.w;
                            s_pixel_lo[column] = (uint16_t)p;
                            src[18 * T] = __uint_as_float((p >> 16) << 8);
                            start_path(src, pi, pj, 0u);
                            have = true;
                            break;
                        }
                    }
                }
                else
                {
                    have = false; // the pixel has all its samples: the column stays out of the next sort
                    if (STABLE)
                    {
                        s_key[column] = 0xFFu;
                    }
                }
            }
        }
    }
    if (!PULL && in_image)
    {
        const float *own = s_state + tid;
        accum[size_t(j) * a.width + i] = make_float4(own[19 * T], own[20 * T], own[21 * T], own[22 * T]);
    }
    if (!PULL && cta_cost && tid == 0)
    {
        cta_cost[strip] = (uint32_t)min((clock64() - t_start) >> 6, 0xFFFFFFFFll);
    }
    uint32_t total = __reduce_add_sync(0xffffffffu, rays);
    uint32_t pixels = PULL ? __reduce_add_sync(0xffffffffu, pixels_done) : __popc(__ballot_sync(0xffffffffu, in_image));
    if (lane == 0 && (pixels || total))
    {
        atomicAdd(counters + 0, (unsigned long long)pixels * a.sample_count);
        atomicAdd(counters + 1, (unsigned long long)total);
    }
    if (COUNT)
    {
        uint32_t v[4] = {cnt.node_visits, cnt.tri_tests, cnt.inst_visits, cnt.sphere_tests};
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            uint32_t sum = __reduce_add_sync(0xffffffffu, v[k]);
            if (lane == 0 && sum)
            {
                atomicAdd(counters + 2 + k, (unsigned long long)sum);
            }
        }
    }
}

// the sorted kernel's dynamic shared memory exceeds the 48 KiB default: opt in once per instantiation
template <int W, int B, int MODE> cudaError_t sorted_allow_smem(int carveout)
{
    static cudaError_t done = [carveout] {
        cudaError_t e = cudaSuccess;
        const int bytes = (int)sorted_smem_bytes<W>();
        // static shared memory next to it: sorted index, (PULL) pixel index, two histogram sets, the total
        const int fixed = W * 32 * 2 * (MODE == 1 ? 2 : 1) + (MODE == 2 ? W * 32 : 0) + 2 * 32 * 4 + 32;
#define SPK_ALLOW(C, N)                                                                                                \
    if (e == cudaSuccess)                                                                                              \
        e = cudaFuncSetAttribute(render_sorted_kernel<C, N, W, B, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                 bytes);                                                                               \
    if (e == cudaSuccess)                                                                                              \
        e = spk::prefer_small_carveout(render_sorted_kernel<C, N, W, B, MODE>, bytes + fixed, B, carveout);
        SPK_ALLOW(true, false)
        SPK_ALLOW(false, false)
        SPK_ALLOW(true, true)
        SPK_ALLOW(false, true)
#undef SPK_ALLOW
        return e;
    }();
    return done;
}

__device__ __forceinline__ float aces(float c)
{
    // ACESFilm (CPURenderer.cpp:554-563): clamp((c(2.51c+0.03))/(c(2.43c+0.59)+0.14), 0, 1)
    float r = (c * (c * 2.51f + 0.03f)) / (c * (c * 2.43f + 0.59f) + 0.14f);
    return fminf(fmaxf(r, 0.f), 1.f);
}

__global__ void probe_libm_kernel(uint32_t op, const float *__restrict__ x, const float *__restrict__ y, uint64_t n,
                                  float *__restrict__ out)
{
    const uint64_t k = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k < n)
    {
        float s, c;
        if (op == 2u)
        {
            out[k] = ref_pow(x[k], y[k]);
        }
        else
        {
            ref_sincos(x[k], &s, &c);
            out[k] = op == 0u ? s : c;
        }
    }
}

__global__ void resolve_kernel(const float4 *__restrict__ accum, uint32_t width, uint32_t height, uint32_t format,
                               float divisor, float exposure, void *__restrict__ out)
{
    uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= width * height)
    {
        return;
    }
    uint32_t i = idx % width, j = idx / width;
    float4 p = accum[idx];
    if (divisor != 1.f)
    {
        p = make_float4(p.x / divisor, p.y / divisor, p.z / divisor, p.w / divisor);
    }
    if (format == SPK_RESOLVE_LINEAR_F32)
    {
        static_cast<float4 *>(out)[idx] = p;
        return;
    }
    float r = aces(p.x * exposure), g = aces(p.y * exposure), b = aces(p.z * exposure);
    size_t o = size_t(height - 1 - j) * width + i; // image.SetPixel(i, H-1-j, ...)
    if (format == SPK_RESOLVE_TONEMAP_F16)
    {
        __half2 lo = __floats2half2_rn(r, g), hi = __floats2half2_rn(b, 1.f);
        uint2 raw;
        raw.x = *reinterpret_cast<uint32_t *>(&lo);
        raw.y = *reinterpret_cast<uint32_t *>(&hi);
        static_cast<uint2 *>(out)[o] = raw;
    }
    else
    {
        // Image2D::SetPixel on an sRGB8 image: LinearToSrgb = pow(0.45), VecToColor = (v*255) truncated
        uchar4 c;
        c.x = (unsigned char)(ref_pow(r, 0.45f) * 255.f);
        c.y = (unsigned char)(ref_pow(g, 0.45f) * 255.f);
        c.z = (unsigned char)(ref_pow(b, 0.45f) * 255.f);
        c.w = 255;
        static_cast<uchar4 *>(out)[o] = c;
    }
}
} // namespace

// Device -> caller's host buffer. Pinned destinations (spk_cuda_host_alloc, cudaHostAlloc, cudaHostRegister) take one
// asynchronous copy at PCIe speed. A pageable destination goes through the context's pinned bounce buffer in 2 MiB
// pieces, the host memcpy of piece k overlapping the DMA of piece k+1 (a plain pageable cudaMemcpy of the 16.6 MB fp16
// frame was ~3 ms of every end-to-end step in round 1).
int copy_to_host(SpkContext *ctx, void *host_out, const void *d_src, size_t bytes)
{
    cudaPointerAttributes attr{};
    const bool pinned = cudaPointerGetAttributes(&attr, host_out) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError(); // an unregistered pointer is not an error here
    if (pinned)
    {
        SPK_CUDA_TRY(cudaMemcpyAsync(host_out, d_src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        return SPK_OK;
    }
    constexpr size_t kPiece = size_t(2) << 20;
    constexpr int kRing = 4;
    if (ctx->h_resolve_capacity < kPiece * kRing)
    {
        cudaFreeHost(ctx->h_resolve);
        ctx->h_resolve = nullptr;
        ctx->h_resolve_capacity = 0;
        SPK_CUDA_TRY(cudaHostAlloc(&ctx->h_resolve, kPiece * kRing, cudaHostAllocDefault));
        ctx->h_resolve_capacity = kPiece * kRing;
        for (int k = 0; k < kRing; k++)
        {
            SPK_CUDA_TRY(cudaEventCreateWithFlags(&ctx->resolve_ev[k], cudaEventDisableTiming));
        }
    }
    const size_t pieces = (bytes + kPiece - 1) / kPiece;
    auto issue = [&](size_t k) -> cudaError_t {
        const size_t at = k * kPiece, n = std::min(kPiece, bytes - at);
        cudaError_t e = cudaMemcpyAsync(static_cast<uint8_t *>(ctx->h_resolve) + (k % kRing) * kPiece,
                                        static_cast<const uint8_t *>(d_src) + at, n, cudaMemcpyDeviceToHost, ctx->stream);
        return e != cudaSuccess ? e : cudaEventRecord(ctx->resolve_ev[k % kRing], ctx->stream);
    };
    for (size_t k = 0; k < pieces && k < size_t(kRing); k++)
    {
        SPK_CUDA_TRY(issue(k));
    }
    for (size_t k = 0; k < pieces; k++)
    {
        SPK_CUDA_TRY(cudaEventSynchronize(ctx->resolve_ev[k % kRing]));
        const size_t at = k * kPiece, n = std::min(kPiece, bytes - at);
        memcpy(static_cast<uint8_t *>(host_out) + at, static_cast<uint8_t *>(ctx->h_resolve) + (k % kRing) * kPiece, n);
        if (k + kRing < pieces)
        {
            SPK_CUDA_TRY(issue(k + kRing));
        }
    }
    return SPK_OK;
}

// ---- cost-ordered strips (owner forms of the sorted megakernel) ----
// d_strip: [cost | cost_sorted | iota | order] of `capacity` entries each + cub scratch. The order belongs to one (strip
// count, form). It is only a schedule - any permutation renders the same image bit for bit - so it is KEPT across camera
// moves and scene uploads: a progressive renderer draws nearly the same view again and again, and after a cut the stale
// order is merely no better than any other. Measured (B200, 1080p): TestScene 256 spp 138.97 -> 130.50 ms, room 943.5 ->
// 933.3 ms (tools/r2_cost_order_ab.py); SPK_OPT_COST_ORDERED_STRIPS = 0 turns it off.
int strip_order_prepare(SpkContext *ctx, uint32_t strips, int form, const uint32_t **order, uint32_t **cost)
{
    if (ctx->strip_capacity < strips)
    {
        cudaFree(ctx->d_strip);
        ctx->d_strip = nullptr;
        ctx->strip_capacity = 0;
        size_t tmp = 0;
        cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                                  (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)strips, 0, 32, ctx->stream);
        ctx->strip_tmp_bytes = tmp;
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_strip, size_t(strips) * 16 + tmp + 256));
        ctx->strip_capacity = strips;
        std::vector<uint32_t> iota(strips);
        for (uint32_t k = 0; k < strips; k++)
        {
            iota[k] = k;
        }
        SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_strip + 2 * size_t(strips), iota.data(), size_t(strips) * 4, cudaMemcpyHostToDevice,
                                     ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        ctx->strip_order_valid = false;
    }
    const size_t cap = ctx->strip_capacity;
    *cost = ctx->d_strip;
    *order = (ctx->strip_order_valid && ctx->strip_order_count == strips && ctx->strip_order_form == form) ? ctx->d_strip + 3 * cap
                                                                                                          : nullptr;
    return SPK_OK;
}

int strip_order_update(SpkContext *ctx, uint32_t strips, int form)
{
    const size_t cap = ctx->strip_capacity;
    size_t tmp = ctx->strip_tmp_bytes;
    SPK_CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(ctx->d_strip + 4 * cap, tmp, ctx->d_strip, ctx->d_strip + cap,
                                                           ctx->d_strip + 2 * cap, ctx->d_strip + 3 * cap, (int)strips, 0, 32,
                                                           ctx->stream));
    ctx->strip_order_valid = true;
    ctx->strip_order_count = strips;
    ctx->strip_order_form = form;
    return SPK_OK;
}

int launch_render(SpkContext *ctx, const SpkRenderParams *p)
{
    RenderArgs a;
    a.width = ctx->width;
    a.height = ctx->height;
    a.first_sample = p->first_sample;
    a.sample_count = p->sample_count;
    a.sample_stride = p->sample_stride ? p->sample_stride : 1u;
    a.max_bounce = p->max_bounce;
    a.output_limit = p->output_limit;
    a.accum_mode = p->accum_mode;
    a.cumulated = ctx->accum_samples;
    a.cumulated_cap = p->cumulated_cap ? p->cumulated_cap : 0xFFFFFFFFu;
    a.debug = (p->flags >> 8) & 3u;
    // spk_cuda_reset_accumulation only marks the accumulator: the megakernels write every pixel, so a fresh launch starts
    // from zero in registers and the 33-133 MB clear never runs; the g-buffer path accumulates in another kernel
    a.fresh = 0u;
    if (ctx->accum_clear_pending)
    {
        if (p->flags & SPK_FLAG_SPATIAL_DENOISE)
        {
            int rc0 = ensure_accum_cleared(ctx);
            if (rc0 != SPK_OK)
            {
                return rc0;
            }
        }
        else
        {
            a.fresh = 1u;
            ctx->accum_clear_pending = false;
        }
    }
    uint32_t tiles = ((a.width + kTileW - 1) / kTileW) * ((a.height + kTileH - 1) / kTileH);
    const DScene &scene = ctx->scene;
    const bool exact = p->flags & SPK_FLAG_EXACT_TRAVERSAL, count = p->flags & SPK_FLAG_COUNTERS;
    // the form of the megakernel: SPK_OPT_MEGAKERNEL_FORM forces one (A/B runs), otherwise the one
    // spk_cuda_choose_tracer measured fastest on this scene (1084 until measured)
    int phased = ctx->opt_mega_form > 0 ? ctx->opt_mega_form : ctx->mega_variant;
    const uint32_t total_px = tiles * 32u;
    if (phased >= 2000 && phased < 3000 && (a.sample_count >= (1u << 17) || total_px >= (1u << 23)))
    {
        phased -= 1000; // the pulling form packs sample : 17 | pixel_hi : 7 | bounce : 8 into one word of a path slot
    }
    // (the owner forms pack sample index : 24 | bounce : 8)
    if (phased >= 1000 && !(p->flags & SPK_FLAG_SPATIAL_DENOISE) && a.max_bounce != 0u && a.debug == 0u &&
        a.max_bounce <= 255u && a.sample_count < (1u << 24))
    {
        static int sms = 0;
        if (sms == 0)
        {
            SPK_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        }
        if (phased >= 2000 && phased < 3000)
        {
            // the pixel counter of the pulling form (counters[7]) starts every launch at zero
            SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_counters + 7, 0, sizeof(unsigned long long), ctx->stream));
        }
        // strips in the order of their cost in the previous launch of this view (owner forms, launches of >= 8 spp)
        const uint32_t strip_warps = uint32_t((phased % 1000) / 10);
        const uint32_t strips = (tiles + strip_warps - 1) / strip_warps;
        const uint32_t *strip_order = nullptr;
        uint32_t *strip_cost = nullptr;
        const bool cost_ordered = ctx->opt_cost_order && (phased < 2000 || phased >= 3000) && a.sample_count >= 8u && strips >= 64u;
        if (cost_ordered)
        {
            int rc1 = strip_order_prepare(ctx, strips, phased, &strip_order, &strip_cost);
            if (rc1 != SPK_OK)
            {
                return rc1;
            }
        }
        // form = 1000 + warps*10 + CTAs/SM: pixels owned by the CTA, one CTA per 16 tiles; 2000 + ...: persistent grid
#define SPK_SORTED_LAUNCH(W, B, MODE, C, N)                                                                            \
    render_sorted_kernel<C, N, W, B, MODE><<<(MODE == 1) ? std::min<uint32_t>((tiles + W - 1) / W, uint32_t(sms) * B)  \
                                                    : (tiles + W - 1) / W,                                             \
                                             W * 32, sorted_smem_bytes<W>(), ctx->stream>>>(                           \
        scene, ctx->camera, a, ctx->d_accum, ctx->d_counters, strip_order, strip_cost)
#define SPK_SORTED_CASE(W, B, MODE)                                                                                    \
    case 1000 * (MODE + 1) + W * 10 + B:                                                                               \
        SPK_CUDA_TRY((sorted_allow_smem<W, B, MODE>(ctx->opt_carveout)));                                              \
        if (!exact && !count)                                                                                          \
            SPK_SORTED_LAUNCH(W, B, MODE, true, false);                                                                \
        else if (exact && !count)                                                                                      \
            SPK_SORTED_LAUNCH(W, B, MODE, false, false);                                                               \
        else if (!exact)                                                                                               \
            SPK_SORTED_LAUNCH(W, B, MODE, true, true);                                                                 \
        else                                                                                                           \
            SPK_SORTED_LAUNCH(W, B, MODE, false, true);                                                                \
        break;
        switch (phased)
        {
            SPK_SORTED_CASE(8, 3, 0)
            SPK_SORTED_CASE(8, 4, 0)
            SPK_SORTED_CASE(16, 2, 0)
            SPK_SORTED_CASE(8, 4, 1)
            SPK_SORTED_CASE(16, 2, 1)
            SPK_SORTED_CASE(8, 4, 2)
            SPK_SORTED_CASE(16, 2, 2)
        default:
            set_error("spk_cuda_render: unknown megakernel form (SPK_OPT_MEGAKERNEL_FORM)");
            return SPK_ERR_INVALID_ARGUMENT;
        }
#undef SPK_SORTED_CASE
#undef SPK_SORTED_LAUNCH
        SPK_CUDA_TRY(cudaGetLastError());
        ctx->stats.kernel_launches++;
        if (cost_ordered)
        {
            return strip_order_update(ctx, strips, phased);
        }
        return SPK_OK;
    }
    if (phased != 1 && !(p->flags & SPK_FLAG_SPATIAL_DENOISE))
    {
        if (phased >= 1000)
        {
            phased = 831; // probes and the zero-bounce case stay on the phased kernel
        }
        // form = warps*100 + min_blocks*10 + second_barrier selects a measured variant; 1 = render_kernel
#define SPK_PHASED_CASE(W, B, S2)                                                                                      \
    case W * 100 + B * 10 + S2:                                                                                        \
        if (!exact && !count)                                                                                          \
            render_phased_kernel<true, false, W, B, S2 != 0>                                                           \
                <<<(tiles + W - 1) / W, W * 32, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_accum, ctx->d_counters); \
        else if (exact && !count)                                                                                      \
            render_phased_kernel<false, false, W, B, S2 != 0>                                                          \
                <<<(tiles + W - 1) / W, W * 32, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_accum, ctx->d_counters); \
        else if (!exact)                                                                                               \
            render_phased_kernel<true, true, W, B, S2 != 0>                                                            \
                <<<(tiles + W - 1) / W, W * 32, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_accum, ctx->d_counters); \
        else                                                                                                           \
            render_phased_kernel<false, true, W, B, S2 != 0>                                                           \
                <<<(tiles + W - 1) / W, W * 32, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_accum, ctx->d_counters); \
        break;
        switch (phased)
        {
            SPK_PHASED_CASE(8, 3, 1)
            SPK_PHASED_CASE(8, 3, 0)
            SPK_PHASED_CASE(16, 2, 1)
            SPK_PHASED_CASE(4, 6, 1)
        default:
            set_error("spk_cuda_render: unknown megakernel form (SPK_OPT_MEGAKERNEL_FORM)");
            return SPK_ERR_INVALID_ARGUMENT;
        }
#undef SPK_PHASED_CASE
        SPK_CUDA_TRY(cudaGetLastError());
        ctx->stats.kernel_launches++;
        return SPK_OK;
    }
    dim3 grid{(tiles + kWarpsPerBlock - 1) / kWarpsPerBlock}, block{kWarpsPerBlock * 32};
    if (p->flags & SPK_FLAG_SPATIAL_DENOISE)
    {
        // the reference's frame: BasePass (one sample into the g-buffer), then DenoisePass (fill + accumulate)
        if (!ctx->d_gbuf)
        {
            SPK_CUDA_TRY(cudaMalloc(&ctx->d_gbuf, size_t(a.width) * a.height * 2 * sizeof(float4)));
        }
        const uint32_t first = a.first_sample, count = a.sample_count;
        a.sample_count = 1;
        for (uint32_t s = 0; s < count; s++)
        {
            a.first_sample = first + s * a.sample_stride;
            if (exact)
                render_kernel<false, false, true><<<grid, block, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_gbuf, ctx->d_counters);
            else
                render_kernel<true, false, true><<<grid, block, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_gbuf, ctx->d_counters);
            SPK_CUDA_TRY(cudaGetLastError());
            ctx->stats.kernel_launches++;
            int rc = launch_denoise_accumulate(ctx, 1u, a.accum_mode, std::min(ctx->accum_samples + s, a.cumulated_cap));
            if (rc != SPK_OK)
            {
                return rc;
            }
        }
        return SPK_OK;
    }
#define SPK_LAUNCH_RENDER(C, N)                                                                                        \
    render_kernel<C, N><<<grid, block, 0, ctx->stream>>>(scene, ctx->camera, a, ctx->d_accum, ctx->d_counters)
    if (exact && count)
        SPK_LAUNCH_RENDER(false, true);
    else if (exact)
        SPK_LAUNCH_RENDER(false, false);
    else if (count)
        SPK_LAUNCH_RENDER(true, true);
    else
        SPK_LAUNCH_RENDER(true, false);
#undef SPK_LAUNCH_RENDER
    SPK_CUDA_TRY(cudaGetLastError());
    ctx->stats.kernel_launches++;
    return SPK_OK;
}
} // namespace spk

using namespace spk;

extern "C"
{
int spk_cuda_render(SpkContext *ctx, const SpkRenderParams *params)
{
    if (!ctx || !params)
    {
        set_error("spk_cuda_render: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_render: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    if (!ctx->has_camera)
    {
        set_error("spk_cuda_render: no camera set");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (params->accum_mode > SPK_ACCUM_RUNNING_AVERAGE || !(params->output_limit > 0.f))
    {
        set_error("spk_cuda_render: bad accum_mode or output_limit");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (ctx->accum_samples != 0 && params->accum_mode != ctx->accum_mode)
    {
        set_error("spk_cuda_render: accum_mode changed without spk_cuda_reset_accumulation");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (params->sample_count == 0)
    {
        return SPK_OK;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
    if ((params->flags & SPK_FLAG_AUTO_TRACER) && ctx->tracer_choice < 0 && ctx->accum_samples == 0 &&
        params->accum_mode == SPK_ACCUM_SUM && params->max_bounce != 0u)
    {
        uint32_t w = 0;
        int rc0 = spk_cuda_choose_tracer(ctx, params->max_bounce, &w, nullptr);
        if (rc0 != SPK_OK)
        {
            return rc0;
        }
    }
    const bool automatic = (params->flags & SPK_FLAG_AUTO_TRACER) && ctx->tracer_choice == 1 &&
                           params->accum_mode == SPK_ACCUM_SUM && ((params->flags >> 8) & 3u) == 0u;
    // a zero-iteration bounce loop has nothing to queue: the megakernel handles it
    const bool wavefront = ((params->flags & SPK_FLAG_WAVEFRONT) || automatic) && params->max_bounce != 0u &&
                           !(params->flags & SPK_FLAG_SPATIAL_DENOISE);
    int rc = wavefront ? launch_render_wavefront(ctx, params) : launch_render(ctx, params);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
    ctx->accum_mode = params->accum_mode;
    ctx->accum_samples += params->sample_count;
    return SPK_OK;
}

int spk_cuda_choose_tracer(SpkContext *ctx, uint32_t max_bounce, uint32_t *wavefront_out, float *ms_out)
{
    if (!ctx || !wavefront_out || max_bounce == 0u)
    {
        set_error("spk_cuda_choose_tracer: null argument or max_bounce 0");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->has_scene || !ctx->has_camera)
    {
        set_error("spk_cuda_choose_tracer: needs a scene and a camera");
        return SPK_ERR_NO_SCENE;
    }
    if (ctx->accum_samples != 0)
    {
        set_error("spk_cuda_choose_tracer: the accumulator is in use (spk_cuda_reset_accumulation first)");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned long long saved[8];
    SPK_CUDA_TRY(cudaMemcpyAsync(saved, ctx->d_counters, sizeof(saved), cudaMemcpyDeviceToHost, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    SpkRenderParams p{};
    p.first_sample = 0;
    p.sample_count = 8;
    p.sample_stride = 1;
    p.max_bounce = max_bounce;
    p.output_limit = 6.f;
    p.accum_mode = SPK_ACCUM_SUM;
    cudaEvent_t e0, e1;
    SPK_CUDA_TRY(cudaEventCreate(&e0));
    SPK_CUDA_TRY(cudaEventCreate(&e1));
    float ms[2] = {3.4e38f, 0.f};
    int rc = SPK_OK;
    // candidates: the sorted megakernel with 8 warps x 4 CTAs/SM and with 16 warps x 2 CTAs/SM (pixels owned by the CTA; the
    // pixel-pulling forms 2162 / 2084 measured 8-15 % slower on TestScene and room and stay as options), then the wavefront tracer
    // (the phased kernel has not won on any scene since the sorted form keeps its path slots in shared memory; it stays
    // as the fallback for probes, zero-bounce and > 255-bounce launches). A later form replaces the incumbent only when
    // it is more than 3 % faster: short timings of near-equal forms are noise.
    static const int kForms[3] = {1084, 1162, -1};
    int best_form = kForms[0];
    for (int which = 0; which < 3 && rc == SPK_OK; which++)
    {
        float best = 3.4e38f;
        for (int rep = 0; rep < 2 && rc == SPK_OK; rep++) // first repetition warms caches and lazy module loading
        {
            if (kForms[which] >= 0)
            {
                ctx->mega_variant = kForms[which];
            }
            cudaEventRecord(e0, ctx->stream);
            rc = kForms[which] < 0 ? launch_render_wavefront(ctx, &p) : launch_render(ctx, &p);
            cudaEventRecord(e1, ctx->stream);
            if (cudaEventSynchronize(e1) != cudaSuccess)
            {
                rc = cuda_fail(cudaGetLastError(), "spk_cuda_choose_tracer");
                break;
            }
            float t = 0.f;
            cudaEventElapsedTime(&t, e0, e1);
            best = t < best ? t : best;
        }
        if (kForms[which] < 0)
        {
            ms[1] = best;
        }
        else if (which == 0 || best < 0.97f * ms[0])
        {
            ms[0] = best;
            best_form = kForms[which];
        }
        if (ctx->opt_verbose)
        {
            fprintf(stderr, "[spk_cuda] choose_tracer: form %d: %.3f ms per %u spp\n", kForms[which], best, p.sample_count);
        }
    }
    ctx->mega_variant = best_form;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    // scratch use only: the accumulator and the counters go back to what they were
    ctx->accum_clear_pending = true;
    SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_counters, saved, sizeof(saved), cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (rc != SPK_OK)
    {
        return rc;
    }
    ctx->tracer_choice = ms[1] < ms[0] ? 1 : 0;
    *wavefront_out = (uint32_t)ctx->tracer_choice;
    if (ms_out)
    {
        ms_out[0] = ms[0];
        ms_out[1] = ms[1];
    }
    return SPK_OK;
}

int spk_cuda_tracer_choice(SpkContext *ctx, int32_t *wavefront_out, uint32_t *megakernel_form_out)
{
    if (!ctx || !wavefront_out || !megakernel_form_out)
    {
        set_error("spk_cuda_tracer_choice: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    *wavefront_out = ctx->tracer_choice;
    *megakernel_form_out = (uint32_t)ctx->mega_variant;
    return SPK_OK;
}

// parity probe: the device's sin / cos / pow on host arrays (tests/test_libm_gpu.py compares them with the host's libm)
int spk_cuda_probe_libm(SpkContext *ctx, uint32_t op, const float *x, const float *y, uint64_t n, float *out)
{
    if (!ctx || !x || !out || op > 2u || (op == 2u && !y) || n > (1ull << 28))
    {
        set_error("spk_cuda_probe_libm: invalid argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (n == 0)
    {
        return SPK_OK;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    float *d = nullptr;
    SPK_CUDA_TRY(cudaMalloc(&d, size_t(n) * 3 * sizeof(float)));
    cudaError_t e = cudaMemcpyAsync(d, x, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess && op == 2u)
    {
        e = cudaMemcpyAsync(d + n, y, n * sizeof(float), cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        probe_libm_kernel<<<unsigned((n + 255) / 256), 256, 0, ctx->stream>>>(op, d, d + n, n, d + 2 * n);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess)
    {
        e = cudaMemcpyAsync(out, d + 2 * n, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        e = cudaStreamSynchronize(ctx->stream);
    }
    cudaFree(d);
    if (e != cudaSuccess)
    {
        set_error(std::string("spk_cuda_probe_libm: ") + cudaGetErrorString(e));
        return SPK_ERR_CUDA;
    }
    return SPK_OK;
}

int spk_cuda_resolve(SpkContext *ctx, uint32_t format, uint32_t total_samples, void *host_out, size_t host_bytes)
{
    if (!ctx || !host_out || format > SPK_RESOLVE_TONEMAP_SRGB8)
    {
        set_error("spk_cuda_resolve: null argument or unknown format");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    size_t pixels = size_t(ctx->width) * ctx->height;
    size_t need = pixels * (format == SPK_RESOLVE_LINEAR_F32 ? 16 : format == SPK_RESOLVE_TONEMAP_F16 ? 8 : 4);
    if (host_bytes < need)
    {
        set_error("spk_cuda_resolve: host buffer too small");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    if (ctx->resolve_capacity < need)
    {
        cudaFree(ctx->d_resolve);
        ctx->d_resolve = nullptr;
        ctx->resolve_capacity = 0;
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_resolve, need));
        ctx->resolve_capacity = need;
    }
    {
        int rc = ensure_accum_cleared(ctx);
        if (rc != SPK_OK)
        {
            return rc;
        }
    }
    uint32_t n = total_samples ? total_samples : ctx->accum_samples;
    float divisor = ctx->accum_mode == SPK_ACCUM_SUM ? (float)(n ? n : 1u) : 1.f;
    float exposure = ctx->has_camera ? ctx->camera.exposure : 1.f;
    unsigned blocks = unsigned((pixels + 255) / 256);
    resolve_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_accum, ctx->width, ctx->height, format, divisor, exposure,
                                                    ctx->d_resolve);
    SPK_CUDA_TRY(cudaGetLastError());
    ctx->stats.kernel_launches++;
    return copy_to_host(ctx, host_out, ctx->d_resolve, need);
}
} // extern "C"
