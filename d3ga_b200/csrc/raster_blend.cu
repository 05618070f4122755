// raster_blend.cu -- rows R6 (alpha blending forward) and R7 (its backward) of SURVEY.md section 8.
//
// Replaces renderCUDA forward / backward inside diff_gaussian_rasterization._C (reached from renderer.py:130-141
// and from autograd when models/trainer.py:186-192 calls loss.backward()).
//
// One CTA per 16x16 tile, one thread per pixel.  The tile's depth-sorted Gaussians are staged through shared
// memory in batches of 256 (id, pixel centre, conic + opacity, colour, 1/depth), front to back in the forward and
// back to front in the backward.  Forward semantics (same order of tests as upstream): skip power > 0, alpha =
// min(.99, o * exp(power)), skip alpha < 1/255, stop the pixel when T * (1 - alpha) < 1e-4, out = C + T * bg.
//
// Backward: every (pixel, Gaussian) pair produces ten partial derivatives.  Instead of ten global atomics per pair
// (upstream) they are reduced in three levels:
//   warp   : butterfly transpose-reduction -- 12 shuffles bring the ten 32-lane sums to ten different lanes
//            (a plain reduction would need 50), skipped entirely when no lane of the warp touches the Gaussian;
//   CTA    : shared-memory accumulators, one row of 12 floats per staged Gaussian;
//   global : three 16-byte vector reductions (REDG.E.ADD.F32x4) per (tile, Gaussian) into the packed accumulator
//            that the preprocess backward (R8) then reads once.
// Algorithmic bytes: forward 40 B per duplicate + 20 B per pixel; backward 40 B read + 40 B gradient RMW per
// duplicate + 20 B per pixel.  Both kernels are FP32/SFU bound rather than HBM bound.
#include "common.cuh"

namespace d3ga {

namespace {

constexpr int BATCH = TILE_PIX;   // Gaussians staged per round

// Pixel footprint of one staged Gaussian: the axis-aligned box outside of which alpha = o * exp(power) is certainly
// below the 1/255 cut (so the exact per-pixel test would skip it anyway).  power >= -tau with tau = ln(255 o) bounds
// d^T C d <= 2 tau; the extents of that ellipse are sqrt(2 tau C^-1_xx), sqrt(2 tau C^-1_yy).  A relative and an
// absolute margin keep the test conservative against rounding; degenerate conics are never culled.
__device__ __forceinline__ float4 footprint_box(float2 c, float4 co) {
    if (co.w < 1.f / 255.f) return make_float4(INFINITY, -INFINITY, INFINITY, -INFINITY);   // can never pass the cut
    const float det = co.x * co.z - co.y * co.y;
    const float tau = __logf(255.f * co.w);
    if (!(det > 0.f) || !(tau >= 0.f)) return make_float4(-INFINITY, INFINITY, -INFINITY, INFINITY);
    const float k = 2.f * tau / det;
    const float hx = sqrtf(k * co.z) * 1.001f + 0.5f, hy = sqrtf(k * co.x) * 1.001f + 0.5f;
    return make_float4(c.x - hx, c.x + hx, c.y - hy, c.y + hy);
}

// Each warp owns an 8x4 pixel block of the 16x16 tile (compact blocks cull better than 16x2 strips).
struct WarpBlock {
    int px, py;
    float x0, x1, y0, y1;
};
__device__ __forceinline__ WarpBlock warp_block(int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    const int bx = blockIdx.x * TILE + (warp & 1) * 8, by = blockIdx.y * TILE + (warp >> 1) * 4;
    WarpBlock w;
    w.px = bx + (lane & 7);
    w.py = by + (lane >> 3);
    w.x0 = (float)bx; w.x1 = (float)(bx + 7); w.y0 = (float)by; w.y1 = (float)(by + 3);
    return w;
}

template <bool INVDEPTH>
__global__ void __launch_bounds__(TILE_PIX) blend_forward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const float2 *__restrict__ xy,
    const float4 *__restrict__ conic_opacity, const float *__restrict__ colors, const float *__restrict__ depths,
    const float *__restrict__ bg, float *__restrict__ final_T, uint32_t *__restrict__ n_contrib,
    float *__restrict__ out_color, float *__restrict__ out_invdepth) {
    __shared__ float2 s_xy[BATCH];
    __shared__ float4 s_co[BATCH];
    __shared__ float4 s_feat[BATCH];   // r, g, b, 1/depth
    __shared__ float4 s_box[BATCH];
    __shared__ uint8_t s_list[TILE_PIX / 32][BATCH];

    const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const WarpBlock wb = warp_block(tid);
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    const uint32_t *point_list = hdr->final_sel ? vals1 : vals0;
    const int rounds = (int)((range.y - range.x + BATCH - 1) / BATCH);
    int todo = (int)(range.y - range.x);

    bool done = !inside;
    float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, invd = 0.f;
    uint32_t last = 0;
    for (int i = 0; i < rounds; i++, todo -= BATCH) {
        if (__syncthreads_count(done) == TILE_PIX) break;
        const uint32_t progress = (uint32_t)i * BATCH + tid;
        if (range.x + progress < range.y) {
            const uint32_t id = point_list[range.x + progress];
            const float2 c = xy[id];
            const float4 co = conic_opacity[id];
            s_xy[tid] = c;
            s_co[tid] = co;
            s_box[tid] = footprint_box(c, co);
            s_feat[tid] = make_float4(colors[3 * (size_t)id], colors[3 * (size_t)id + 1], colors[3 * (size_t)id + 2],
                                      INVDEPTH ? __fdividef(1.f, depths[id]) : 0.f);
        }
        __syncthreads();
        if (__all_sync(0xffffffffu, done)) continue;
        const int n = min(BATCH, todo);
        // this warp's depth-ordered list of the staged Gaussians whose footprint reaches its pixel block
        int cnt = 0;
#pragma unroll
        for (int r = 0; r < BATCH / 32; r++) {
            const int g = r * 32 + (int)lane;
            const float4 b = s_box[g];
            const bool hit = g < n && b.y >= wb.x0 && b.x <= wb.x1 && b.w >= wb.y0 && b.z <= wb.y1;
            const uint32_t m = __ballot_sync(0xffffffffu, hit);
            if (hit) s_list[warp][cnt + __popc(m & lanemask_lt())] = (uint8_t)g;
            cnt += __popc(m);
        }
        __syncwarp();
        const uint32_t base = (uint32_t)i * BATCH;
        for (int k = 0; k < cnt; k++) {
            if (__all_sync(0xffffffffu, done)) break;
            if (done) continue;
            const int j = s_list[warp][k];
            const float2 c = s_xy[j];
            const float dx = c.x - pxf, dy = c.y - pyf;
            const float4 co = s_co[j];
            const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
            if (power > 0.f) continue;
            const float alpha = fminf(0.99f, co.w * __expf(power));
            if (alpha < 1.f / 255.f) continue;
            const float test_T = T * (1.f - alpha);
            if (test_T < 0.0001f) { done = true; continue; }
            const float4 f = s_feat[j];
            const float w = alpha * T;
            C0 += f.x * w; C1 += f.y * w; C2 += f.z * w;
            if (INVDEPTH) invd += f.w * w;
            T = test_T;
            last = base + (uint32_t)j + 1u;
        }
    }
    if (inside) {
        const size_t pid = (size_t)py * W + px, plane = (size_t)H * W;
        final_T[pid] = T;
        n_contrib[pid] = last;
        out_color[pid] = C0 + T * bg[0];
        out_color[plane + pid] = C1 + T * bg[1];
        out_color[2 * plane + pid] = C2 + T * bg[2];
        if (INVDEPTH) out_invdepth[pid] = invd;
    }
}

// Ten values per lane, summed over the 32 lanes, end up in ten different lanes after 12 shuffles.
//   stage xor16: 10 -> 5 values per lane   stage xor8: 5 -> 3   stage xor4: 3 -> 2   stage xor2: 2 -> 1   stage xor1
// slot_of_lane() tells which of the ten sums a lane holds afterwards (-1: none; odd lanes duplicate the even ones).
__device__ __forceinline__ int slot_of_lane(uint32_t lane) {
    if (lane & 1u) return -1;
    int base = (lane & 16u) ? 5 : 0, n = 5;
    if (lane & 8u) { base += 3; n = 2; } else { n = 3; }
    if (lane & 4u) { base += 2; n -= 2; } else { n = min(n, 2); }
    if (lane & 2u) { base += 1; n -= 1; } else { n = min(n, 1); }
    return n >= 1 ? base : -1;
}
__device__ __forceinline__ float butterfly10(const float (&v)[10], uint32_t lane) {
    const bool b4 = lane & 16u, b3 = lane & 8u, b2 = lane & 4u, b1 = lane & 2u;
    float a[5];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        const float send = b4 ? v[k] : v[k + 5], keep = b4 ? v[k + 5] : v[k];
        a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
    float b[3];
    {
        const float r0 = __shfl_xor_sync(0xffffffffu, b3 ? a[0] : a[3], 8);
        const float r1 = __shfl_xor_sync(0xffffffffu, b3 ? a[1] : a[4], 8);
        const float r2 = __shfl_xor_sync(0xffffffffu, b3 ? a[2] : 0.f, 8);
        b[0] = (b3 ? a[3] : a[0]) + r0;
        b[1] = (b3 ? a[4] : a[1]) + r1;
        b[2] = b3 ? 0.f : a[2] + r2;
    }
    float c[2];
    {
        const float r0 = __shfl_xor_sync(0xffffffffu, b2 ? b[0] : b[2], 4);
        const float r1 = __shfl_xor_sync(0xffffffffu, b2 ? b[1] : 0.f, 4);
        c[0] = (b2 ? b[2] : b[0]) + r0;
        c[1] = b2 ? 0.f : b[1] + r1;
    }
    float d = (b1 ? c[1] : c[0]) + __shfl_xor_sync(0xffffffffu, b1 ? c[0] : c[1], 2);
    d += __shfl_xor_sync(0xffffffffu, d, 1);
    return d;
}

template <bool INVDEPTH>
__global__ void __launch_bounds__(TILE_PIX) blend_backward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const float *__restrict__ bg, const float2 *__restrict__ xy,
    const float4 *__restrict__ conic_opacity, const float *__restrict__ colors, const float *__restrict__ depths,
    const float *__restrict__ final_T, const uint32_t *__restrict__ n_contrib, const float *__restrict__ dL_dpix,
    const float *__restrict__ dL_dinvdepth_pix, float *__restrict__ grad_acc) {
    __shared__ float2 s_xy[BATCH];
    __shared__ float4 s_co[BATCH];
    __shared__ float4 s_feat[BATCH];
    __shared__ float4 s_box[BATCH];
    __shared__ uint32_t s_id[BATCH];
    __shared__ uint8_t s_list[TILE_PIX / 32][BATCH];
    __shared__ __align__(16) float s_acc[BATCH * GRAD_ACC_FLOATS];
    __shared__ uint32_t s_max[TILE_PIX / 32];

    const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const WarpBlock wb = warp_block(tid);
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    if (range.y <= range.x) return;
    const uint32_t *point_list = hdr->final_sel ? vals1 : vals0;
    const size_t pid = (size_t)py * W + px, plane = (size_t)H * W;

    const float T_final = inside ? final_T[pid] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pid] : 0u;
    // Only the first n_proc Gaussians of the tile's list reached any pixel in the forward: nothing behind them is
    // staged at all (the forward stops a pixel at T < 1e-4, typically after a small fraction of a deep list).
    const uint32_t warp_last = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) s_max[warp] = warp_last;
    for (int k = tid; k < BATCH * GRAD_ACC_FLOATS; k += TILE_PIX) s_acc[k] = 0.f;
    __syncthreads();
    uint32_t n_proc = 0;
#pragma unroll
    for (int w = 0; w < TILE_PIX / 32; w++) n_proc = max(n_proc, s_max[w]);
    if (n_proc == 0) return;
    const int rounds = (int)((n_proc + BATCH - 1) / BATCH);

    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acci = 0.f;
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, lci = 0.f, last_alpha = 0.f;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f, gi = 0.f;
    if (inside) {
        g0 = dL_dpix[pid]; g1 = dL_dpix[plane + pid]; g2 = dL_dpix[2 * plane + pid];
        if (INVDEPTH) gi = dL_dinvdepth_pix[pid];
    }
    const float bg_dot = bg[0] * g0 + bg[1] * g1 + bg[2] * g2;
    const float ddelx_dx = 0.5f * (float)W, ddely_dy = 0.5f * (float)H;
    const int slot = slot_of_lane(lane);

    for (int i = 0; i < rounds; i++) {
        __syncthreads();
        // staged slot t holds the Gaussian at forward position q = n_proc - 1 - (i * BATCH + t): back to front
        const uint32_t progress = (uint32_t)i * BATCH + tid;
        if (progress < n_proc) {
            const uint32_t id = point_list[range.x + (n_proc - 1u - progress)];
            const float2 c = xy[id];
            const float4 co = conic_opacity[id];
            s_id[tid] = id;
            s_xy[tid] = c;
            s_co[tid] = co;
            s_box[tid] = footprint_box(c, co);
            s_feat[tid] = make_float4(colors[3 * (size_t)id], colors[3 * (size_t)id + 1], colors[3 * (size_t)id + 2],
                                      INVDEPTH ? __fdividef(1.f, depths[id]) : 0.f);
        }
        __syncthreads();
        const int n = (int)min((uint32_t)BATCH, n_proc - (uint32_t)i * BATCH);
        const uint32_t q_first = n_proc - 1u - (uint32_t)i * BATCH;   // forward position of staged slot 0
        int cnt = 0;
#pragma unroll
        for (int r = 0; r < BATCH / 32; r++) {
            const int g = r * 32 + (int)lane;
            const float4 b = s_box[g];
            const bool hit = g < n && (q_first - (uint32_t)g) < warp_last && b.y >= wb.x0 && b.x <= wb.x1 &&
                             b.w >= wb.y0 && b.z <= wb.y1;
            const uint32_t m = __ballot_sync(0xffffffffu, hit);
            if (hit) s_list[warp][cnt + __popc(m & lanemask_lt())] = (uint8_t)g;
            cnt += __popc(m);
        }
        __syncwarp();
        for (int k = 0; k < cnt; k++) {
            const int j = s_list[warp][k];
            const uint32_t q = q_first - (uint32_t)j;
            float v[10];
#pragma unroll
            for (int e = 0; e < 10; e++) v[e] = 0.f;
            bool hit = false;
            if (q < last_contributor) {
                const float2 c = s_xy[j];
                const float dx = c.x - pxf, dy = c.y - pyf;
                const float4 co = s_co[j];
                const float power = -0.5f * (co.x * dx * dx + co.z * dy * dy) - co.y * dx * dy;
                if (power <= 0.f) {
                    const float G = __expf(power);
                    const float alpha = fminf(0.99f, co.w * G);
                    if (alpha >= 1.f / 255.f) {
                        hit = true;
                        const float inv_1ma = __fdividef(1.f, 1.f - alpha);
                        T = T * inv_1ma;
                        const float dchannel_dcolor = alpha * T;
                        const float4 f = s_feat[j];
                        acc0 = last_alpha * lc0 + (1.f - last_alpha) * acc0; lc0 = f.x;
                        acc1 = last_alpha * lc1 + (1.f - last_alpha) * acc1; lc1 = f.y;
                        acc2 = last_alpha * lc2 + (1.f - last_alpha) * acc2; lc2 = f.z;
                        float dL_dalpha = (f.x - acc0) * g0 + (f.y - acc1) * g1 + (f.z - acc2) * g2;
                        v[6] = dchannel_dcolor * g0; v[7] = dchannel_dcolor * g1; v[8] = dchannel_dcolor * g2;
                        if (INVDEPTH) {
                            acci = last_alpha * lci + (1.f - last_alpha) * acci; lci = f.w;
                            dL_dalpha += (f.w - acci) * gi;
                            v[9] = dchannel_dcolor * gi;
                        }
                        dL_dalpha *= T;
                        last_alpha = alpha;
                        dL_dalpha += (-T_final * inv_1ma) * bg_dot;
                        const float dL_dG = co.w * dL_dalpha;
                        const float gdx = G * dx, gdy = G * dy;
                        const float dG_ddelx = -gdx * co.x - gdy * co.y;
                        const float dG_ddely = -gdy * co.z - gdx * co.y;
                        v[0] = dL_dG * dG_ddelx * ddelx_dx;
                        v[1] = dL_dG * dG_ddely * ddely_dy;
                        v[2] = -0.5f * gdx * dx * dL_dG;
                        v[3] = -0.5f * gdx * dy * dL_dG;
                        v[4] = -0.5f * gdy * dy * dL_dG;
                        v[5] = G * dL_dalpha;
                    }
                }
            }
            if (!__any_sync(0xffffffffu, hit)) continue;
            const float sum = butterfly10(v, lane);
            if (slot >= 0 && sum != 0.f) atomicAdd(&s_acc[j * GRAD_ACC_FLOATS + slot], sum);
        }
        __syncthreads();
        // flush this batch: one row of 12 floats per staged Gaussian -> three vector reductions
        if (tid < n) {
            float4 *row = reinterpret_cast<float4 *>(s_acc + tid * GRAD_ACC_FLOATS);
            const float4 r0 = row[0], r1 = row[1], r2 = row[2];
            float *dst = grad_acc + (size_t)s_id[tid] * GRAD_ACC_FLOATS;
            if (r0.x != 0.f || r0.y != 0.f || r0.z != 0.f || r0.w != 0.f) red_add_v4(dst, r0.x, r0.y, r0.z, r0.w);
            if (r1.x != 0.f || r1.y != 0.f || r1.z != 0.f || r1.w != 0.f) red_add_v4(dst + 4, r1.x, r1.y, r1.z, r1.w);
            if (r2.x != 0.f || r2.y != 0.f) red_add_v2(dst + 8, r2.x, r2.y);
            row[0] = make_float4(0.f, 0.f, 0.f, 0.f);
            row[1] = make_float4(0.f, 0.f, 0.f, 0.f);
            row[2] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}

}  // namespace

int launch_blend_forward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                         const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, uint8_t *image,
                         const ImgLayout &IL, float *out_color, float *out_invdepth, cudaStream_t stream) {
    const dim3 grid((W + TILE - 1) / TILE, (H + TILE - 1) / TILE), block(TILE_PIX);
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const float *depth = reinterpret_cast<const float *>(geom + GL.depth);
    float *final_T = reinterpret_cast<float *>(image + IL.final_T);
    uint32_t *n_contrib = reinterpret_cast<uint32_t *>(image + IL.n_contrib);
    ProfScope prof(K_BLEND_FWD, stream);
    if (out_invdepth)
        blend_forward_kernel<true><<<grid, block, 0, stream>>>(hdr, ranges, v0, v1, W, H, xy, co, colors, depth,
                                                               background, final_T, n_contrib, out_color, out_invdepth);
    else
        blend_forward_kernel<false><<<grid, block, 0, stream>>>(hdr, ranges, v0, v1, W, H, xy, co, colors, depth,
                                                                background, final_T, n_contrib, out_color, nullptr);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

int launch_blend_backward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                          const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, const uint8_t *image,
                          const ImgLayout &IL, const float *dL_dpix, const float *dL_dinvdepth_pix, float *grad_acc,
                          cudaStream_t stream) {
    const dim3 grid((W + TILE - 1) / TILE, (H + TILE - 1) / TILE), block(TILE_PIX);
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const float *depth = reinterpret_cast<const float *>(geom + GL.depth);
    const float *final_T = reinterpret_cast<const float *>(image + IL.final_T);
    const uint32_t *n_contrib = reinterpret_cast<const uint32_t *>(image + IL.n_contrib);
    ProfScope prof(K_BLEND_BWD, stream);
    if (dL_dinvdepth_pix)
        blend_backward_kernel<true><<<grid, block, 0, stream>>>(hdr, ranges, v0, v1, W, H, background, xy, co, colors,
                                                                depth, final_T, n_contrib, dL_dpix, dL_dinvdepth_pix,
                                                                grad_acc);
    else
        blend_backward_kernel<false><<<grid, block, 0, stream>>>(hdr, ranges, v0, v1, W, H, background, xy, co, colors,
                                                                 depth, final_T, n_contrib, dL_dpix, nullptr, grad_acc);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
