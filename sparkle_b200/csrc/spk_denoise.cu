/*
 * spk_denoise.cu — CPURenderer::DenoisePass (libraries/source/renderer/renderer/CPURenderer.cpp:506-552) for one
 * sample per pixel held in a g-buffer: the spatial fill of SpatialDenoisePixel (:452-504) and the temporal accumulate
 * (:530-544), fused into one pass over the image (2 x 16 B read of the pixel, neighbour reads only around invalid
 * pixels, 16 B read-modify-write of the accumulator: HBM bound, ~64 B per pixel).
 *
 * The reference SCATTERS: every valid non-sky pixel p gives its colour to the first of its 8 neighbours (in the
 * Directions order) that is invalid and whose normal is within 0.99 of p's, then stops. Several p can pick the same
 * target; run on one worker the rows go in order, so the last p in row-major order wins (with several workers the
 * reference's winner is a race). This kernel GATHERS the same result: an invalid pixel q looks at the 8 pixels that
 * could target it, keeps those whose first acceptable target is q, and takes the last of them in row-major order.
 */
#include "spk_context.h"

namespace spk
{
namespace
{
__constant__ int c_dirs[8][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}, {-1, -1}, {1, -1}, {-1, 1}, {1, 1}};

__device__ __forceinline__ bool acceptable(const float4 *__restrict__ color, const float4 *__restrict__ normal, int w,
                                           int h, int ti, int tj, float4 n_src)
{
    // the tests SpatialDenoisePixel applies to a neighbour (:474-497)
    if (ti < 0 || ti >= w || tj < 0 || tj >= h)
    {
        return false;
    }
    size_t t = size_t(tj) * w + ti;
    if (color[t].w > 0.f)
    {
        return false;
    }
    float4 nn = normal[t];
    float similarity = (nn.x * n_src.x + nn.y * n_src.y) + nn.z * n_src.z; // neighbour_normal.dot(world_normal)
    return !(similarity < 0.99f);
}

__global__ void denoise_accumulate_kernel(const float4 *__restrict__ gbuf, float4 *__restrict__ accum, uint32_t width,
                                          uint32_t height, uint32_t spatial, uint32_t accum_mode, uint32_t cumulated)
{
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= width * height)
    {
        return;
    }
    const float4 *__restrict__ color = gbuf;
    const float4 *__restrict__ normal = gbuf + size_t(width) * height;
    const int w = int(width), h = int(height);
    const int qi = int(idx % width), qj = int(idx / width);
    float4 px = color[idx];
    if (spatial && !(px.w > 0.f))
    {
        long best = -1;
        for (int k = 0; k < 8; k++)
        {
            // p gives to q through direction k: p = q - dir[k]
            const int pi = qi - c_dirs[k][0], pj = qj - c_dirs[k][1];
            if (pi < 0 || pi >= w || pj < 0 || pj >= h)
            {
                continue;
            }
            const size_t p = size_t(pj) * w + pi;
            const float4 pc = color[p], pn = normal[p];
            // only a valid non-sky intersection may be used as reference (:463-467)
            if (!(pc.w > 0.f) || (pn.x == 0.f && pn.y == 0.f && pn.z == 0.f))
            {
                continue;
            }
            bool first = acceptable(color, normal, w, h, qi, qj, pn);
            for (int e = 0; e < k && first; e++)
            {
                first = !acceptable(color, normal, w, h, pi + c_dirs[e][0], pj + c_dirs[e][1], pn);
            }
            if (first && long(p) > best)
            {
                best = long(p);
            }
        }
        if (best >= 0)
        {
            float4 c = color[best];
            px = make_float4(c.x, c.y, c.z, 1.f);
        }
    }
    float4 acc = accum[idx];
    if (accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
    {
        // utilities::Lerp(new, acc, N / (N + 1)) = new + (acc - new) * ma  (Utilities.h:119-122)
        float ma = (float)cumulated / (float)(cumulated + 1u);
        acc = make_float4(px.x + (acc.x - px.x) * ma, px.y + (acc.y - px.y) * ma, px.z + (acc.z - px.z) * ma,
                          px.w + (acc.w - px.w) * ma);
    }
    else
    {
        acc = make_float4(acc.x + px.x, acc.y + px.y, acc.z + px.z, acc.w + px.w);
    }
    accum[idx] = acc;
}
} // namespace

int launch_denoise_accumulate(SpkContext *ctx, uint32_t spatial, uint32_t accum_mode, uint32_t cumulated)
{
    const size_t pixels = size_t(ctx->width) * ctx->height;
    unsigned blocks = unsigned((pixels + 255) / 256);
    denoise_accumulate_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->d_gbuf, ctx->d_accum, ctx->width, ctx->height, spatial,
                                                             accum_mode, cumulated);
    SPK_CUDA_TRY(cudaGetLastError());
    ctx->stats.kernel_launches++;
    return SPK_OK;
}
} // namespace spk
