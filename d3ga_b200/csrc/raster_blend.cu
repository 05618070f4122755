// raster_blend.cu -- rows R6 (alpha blending forward) and R7 (its backward) of SURVEY.md section 8.
//
// Replaces renderCUDA forward / backward inside diff_gaussian_rasterization._C (reached from renderer.py:130-141
// and from autograd when models/trainer.py:186-192 calls loss.backward()).
//
// One CTA per 16x16 tile (longest list first), one thread per pixel, one warp per 8x4 pixel block.  Each warp walks the
// tile's depth-sorted list ON ITS OWN -- 32 Gaussians per step, one per lane, gathered a step ahead in registers (ids two
// steps ahead), tested exactly against the warp's pixel block (footprint_hit), and only the survivors are parked
// (compacted, in list order) in a per-warp ring in shared memory from which the blend loop takes them in groups.  There
// is no CTA barrier in the loop: with batches staged by the whole CTA a third of all warp-stall samples sat on that
// barrier, because per-warp culling gives the eight warps of a tile very different amounts of work.  A warp whose pixels
// are all finished simply stops.  The price is that the eight warps each read the tile's list (they run close enough
// together that seven of them hit in L1) and each evaluate the footprint test.  Front to back in the forward, back to
// front over the prefix its own pixels reached in the backward.
// Forward semantics (same order of tests as upstream): skip power > 0, alpha = min(.99, o * exp(power)), skip
// alpha < 1/255, stop the pixel when T * (1 - alpha) < 1e-4, out = C + T * bg.
//
// Backward: every (pixel, Gaussian) pair owes the per-Gaussian accumulator nine or ten partial sums: the six raw moments
// of p = G dL/dalpha about the Gaussian's 2-D centre (sum p dx, p dy, p dx^2, p dx dy, p dy^2, p -- the preprocess backward
// turns them into dL/dmean2D, dL/dconic, dL/dopacity with the per-Gaussian conic, so the pixel loop never touches the
// conic again), the three colour derivatives and the inverse-depth one.  Instead of ten global atomics per pair (upstream)
// the pixel loop parks two scalars per pair in shared memory and a per-warp "turn-around" every 16 Gaussians forms the
// sums and issues vector reductions -- see the comment above backward_turn_around.
// dL/dalpha uses upstream's recurrence contracted with dL/dpixel BEFORE the loop instead of after (one scalar "behind"
// = sum over the contributors behind of (c . g) alpha T instead of a running colour per channel).
// Algorithmic bytes: forward 40 B per duplicate + 20 B per pixel; backward 40 B read + 40 B gradient RMW per
// duplicate + 20 B per pixel.  Both kernels are issue / latency bound rather than HBM bound.
// Measured and dropped: persistent CTAs with a per-CTA tile queue or a per-CTA block dispenser (the warps of a tile drift
// apart, the list no longer hits in L1: 273 -> 333-386 us backward), 2 or 4 warps per CTA, 4 CTAs/SM in the backward.
#include "common.cuh"

namespace d3ga {

namespace {

// Conic pre-scaled for the inner loop: log2(G) = A dx^2 + B dx dy + C dy^2 with A = -0.5 log2(e) conic.xx, B = -log2(e)
// conic.xy, C = -0.5 log2(e) conic.yy, so G = exp2(.) costs one MUFU and no extra multiply.
constexpr float LOG2E = 1.4426950408889634f;
__device__ __forceinline__ float4 prescale_conic(float4 co) {
    return make_float4(-0.5f * LOG2E * co.x, -LOG2E * co.y, -0.5f * LOG2E * co.z, co.w);
}
__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Each warp owns an 8x4 pixel block of the 16x16 tile (compact blocks cull better than 16x2 strips).
struct WarpBlock {
    int px, py;
    float x0, x1, y0, y1;
};
__device__ __forceinline__ WarpBlock warp_block(int warp, int lane, int tile_x, int tile_y) {
    const int bx = tile_x * TILE + (warp & 1) * 8, by = tile_y * TILE + (warp >> 1) * 4;
    WarpBlock w;
    w.px = bx + (lane & 7);
    w.py = by + (lane >> 3);
    w.x0 = (float)bx; w.x1 = (float)(bx + 7); w.y0 = (float)by; w.y1 = (float)(by + 3);
    return w;
}
__device__ __forceinline__ WarpBlock warp_block(int tid, int tile_x, int tile_y) {
    return warp_block(tid >> 5, tid & 31, tile_x, tile_y);
}

constexpr int RING = 64;                     // per-warp ring of staged survivors (>= group - 1 + 32)
constexpr int FWD_WARPS = TILE_PIX / 32;    // (splitting a tile over 2 or 4 smaller CTAs was measured: no gain)
#ifndef FWD_MIN_CTAS
#define FWD_MIN_CTAS 3                       // 80 registers, no spills: 165 us; 4 CTAs/SM (64 registers, spills): 181 us
#endif
#ifndef BWD_MIN_CTAS
#define BWD_MIN_CTAS 3
#endif
#ifdef D3GA_BLEND_STATS
// development counters: [0..3] forward (warp-steps, survivors, groups, passing pixel-Gaussian pairs), [4..7] backward
__device__ unsigned long long g_blend_stats[8];
__device__ unsigned long long *g_blend_trace = nullptr;   // [tiles * 8][4]: start ns, end ns, (smid << 32 | steps), survivors
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#ifdef D3GA_BLEND_TRACE_ONLY   // the counters cost a global atomic per step; the trace alone does not perturb the timing
#define STAT_ADD(i, v) do { } while (0)
#else
#define STAT_ADD(i, v) do { const unsigned long long _t = (unsigned long long)(v); if (_t) atomicAdd(&g_blend_stats[i], _t); } while (0)
#endif
#else
#define STAT_ADD(i, v) do { } while (0)
#endif

// Exact version of the footprint test for one warp's pixel block: does the ellipse d^T C d <= 2 tau (outside of which
// alpha < 1/255: power >= -tau with tau = ln(255 o) bounds d^T C d <= 2 tau) reach any pixel centre of the block?  The quadratic is convex with its minimum at
// the Gaussian centre, so when the centre lies outside the block its minimum over the block sits on the edge(s) facing
// the centre; on an edge it is a 1-D quadratic, minimised in closed form and clamped.  A relative and an absolute
// margin keep the test conservative against rounding (a false "hit" only costs time, the per-pixel test decides);
// degenerate conics are never culled.  Rejects the diagonal, elongated splats the bounding box lets through.
__device__ __forceinline__ bool footprint_hit(float2 c, float4 co, const WarpBlock &wb) {
    if (co.w < 1.f / 255.f) return false;   // can never pass the cut
    const float det = co.x * co.z - co.y * co.y;
    const float tau = __logf(255.f * co.w);
    if (!(det > 0.f) || !(tau >= 0.f) || !(co.x > 0.f) || !(co.z > 0.f)) return true;
    const float lim = 2.f * tau * 1.002f + 0.01f;
    const float xl = wb.x0 - c.x, xh = wb.x1 - c.x, yl = wb.y0 - c.y, yh = wb.y1 - c.y;
    const bool in_x = xl <= 0.f && xh >= 0.f, in_y = yl <= 0.f && yh >= 0.f;
    if (in_x && in_y) return true;
    float qmin = INFINITY;
    if (!in_x) {
        const float dx = xl > 0.f ? xl : xh;
        const float dy = fminf(fmaxf(__fdividef(-co.y * dx, co.z), yl), yh);
        qmin = co.x * dx * dx + 2.f * co.y * dx * dy + co.z * dy * dy;
    }
    if (!in_y) {
        const float dy = yl > 0.f ? yl : yh;
        const float dx = fminf(fmaxf(__fdividef(-co.y * dy, co.x), xl), xh);
        qmin = fminf(qmin, co.x * dx * dx + 2.f * co.y * dx * dy + co.z * dy * dy);
    }
    return qmin <= lim;
}

template <bool INVDEPTH>
__global__ void __launch_bounds__(FWD_WARPS * 32, FWD_MIN_CTAS) blend_forward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const float2 *__restrict__ xy,
    const float4 *__restrict__ conic_opacity, const float *__restrict__ colors, const float *__restrict__ depths,
    const float *__restrict__ bg, float *__restrict__ final_T, uint32_t *__restrict__ n_contrib,
    float *__restrict__ out_color, float *__restrict__ out_invdepth, RasterHeader *__restrict__ hdr_rw, int grid_x,
    int num_tiles, const uint32_t *__restrict__ order_fwd, uint32_t *__restrict__ tile_nproc,
    uint32_t *__restrict__ order_bwd) {
    __shared__ float4 s_a[FWD_WARPS][RING];      // x, y, list position + 1 (bits), unused
    __shared__ float4 s_co[FWD_WARPS][RING];     // pre-scaled conic, opacity
    __shared__ float4 s_feat[FWD_WARPS][RING];   // r, g, b, 1/depth
    __shared__ uint32_t s_bins[257];
    __shared__ uint32_t s_last;

    // The blend loop below is branch-free: entries of a partly filled last group are multiplied by a weight of zero
    // rather than skipped, so a ring slot that was never written must not hold a NaN or Inf bit pattern.
    for (int i = threadIdx.x; i < FWD_WARPS * RING; i += FWD_WARPS * 32) (&s_feat[0][0])[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    constexpr uint32_t SPLIT = 8 / FWD_WARPS;          // CTAs per tile
    const uint32_t tile = order_fwd[blockIdx.x / SPLIT];   // longest Gaussian list first
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const WarpBlock wb = warp_block((int)((blockIdx.x % SPLIT) * FWD_WARPS + warp), (int)lane,
                                    (int)(tile % (uint32_t)grid_x), (int)(tile / (uint32_t)grid_x));
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    const uint32_t *point_list = (hdr->final_sel ? vals1 : vals0) + range.x;
    const uint32_t count = range.y - range.x;
    const uint32_t nchunks = (count + 31u) / 32u;

    bool done = !inside;
    float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, invd = 0.f;
    uint32_t last = 0;
    float4(*ra)[RING] = s_a, (*rco)[RING] = s_co, (*rfeat)[RING] = s_feat;
#ifdef D3GA_BLEND_STATS
    const unsigned long long tr_t0 = global_ns();
    unsigned tr_steps = 0, tr_surv = 0;
#endif

    float2 p_xy = make_float2(0.f, 0.f);
    float4 p_co = make_float4(0.f, 0.f, 0.f, 0.f), p_feat = make_float4(0.f, 0.f, 0.f, 0.f);
    auto gather = [&](uint32_t id) {
        p_xy = xy[id];
        p_co = conic_opacity[id];
        p_feat = make_float4(colors[3 * (size_t)id], colors[3 * (size_t)id + 1], colors[3 * (size_t)id + 2],
                             INVDEPTH ? depths[id] : 1.f   /* raw depth: dividing here would wait for the load */);
    };
    // ids run two steps ahead of the blend, the Gaussians themselves one step
    uint32_t id_next = 0;
    if (lane < count) gather(point_list[lane]);
    if (32u + lane < count) id_next = point_list[32u + lane];
    uint32_t head = 0, tail = 0;
    const bool warp_active = !__all_sync(0xffffffffu, done);
    for (uint32_t c = 0; c < nchunks && warp_active; c++) {
        const uint32_t pos = c * 32u + lane;
        bool hit = false;
        if (pos < count) hit = footprint_hit(p_xy, p_co, wb);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) {
            const uint32_t slot = (tail + __popc(m & lanemask_lt())) & (RING - 1);
            ra[warp][slot] = make_float4(p_xy.x, p_xy.y, __uint_as_float(pos + 1u), 0.f);
            rco[warp][slot] = prescale_conic(p_co);
            rfeat[warp][slot] = make_float4(p_feat.x, p_feat.y, p_feat.z, INVDEPTH ? __fdividef(1.f, p_feat.w) : 0.f);
        }
        tail += __popc(m);
        if (lane == 0) { STAT_ADD(0, 1); STAT_ADD(1, __popc(m)); }
#ifdef D3GA_BLEND_STATS
        tr_steps++; tr_surv += __popc(m);
#endif
        if (pos + 32u < count) gather(id_next);                      // in flight while the ring is blended
        if (pos + 64u < count) id_next = point_list[pos + 64u];
        __syncwarp();
        const bool flush = c + 1u == nchunks;
        bool all_done = false;
        // four ring entries at a time (head stays a multiple of four, so a group never wraps): their alphas are
        // independent; the transmittance recurrence that follows is branch-free
        while (tail - head >= 4u || (flush && tail != head)) {
            const uint32_t avail = min(4u, tail - head);
            const uint32_t base = head & (RING - 1);
            const float4 *ga = &ra[warp][base], *gco = &rco[warp][base], *gf = &rfeat[warp][base];
            float a[4];
            uint32_t posv[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 e = ga[u];
                const float dx = e.x - pxf, dy = e.y - pyf;
                const float4 co = gco[u];   // pre-scaled: log2 domain
                const float power = dx * (co.x * dx + co.y * dy) + co.z * dy * dy;
                const float alpha = fminf(0.99f, co.w * fast_exp2(power));
                a[u] = ((uint32_t)u < avail && !(power > 0.f) && alpha >= 1.f / 255.f) ? alpha : 0.f;
                posv[u] = __float_as_uint(e.z);
            }
            if (lane == 0) STAT_ADD(2, 1);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 f = gf[u];
                const float test_T = T * (1.f - a[u]);
                const bool live = !done && a[u] != 0.f;
                const bool stop = live && test_T < 0.0001f;
                const bool upd = live && !stop;
                done = done || stop;
                const float w = upd ? a[u] * T : 0.f;
                C0 += f.x * w; C1 += f.y * w; C2 += f.z * w;
                if (INVDEPTH) invd += f.w * w;
                T = upd ? test_T : T;
                last = upd ? posv[u] : last;
            }
            head += avail;
            if (__all_sync(0xffffffffu, done)) { all_done = true; break; }
        }
        if (all_done) break;
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
#ifdef D3GA_BLEND_STATS
    if (g_blend_trace && lane == 0) {
        unsigned sm;
        asm("mov.u32 %0, %%smid;" : "=r"(sm));
        unsigned long long *o = g_blend_trace + 4 * ((size_t)blockIdx.x * FWD_WARPS + warp);
        o[0] = tr_t0; o[1] = global_ns(); o[2] = ((unsigned long long)sm << 32) | tr_steps; o[3] = ((unsigned long long)count << 32) | tr_surv;
    }
#endif
    // deepest contributor of the tile = how far the backward has to go; the last CTA orders the backward's tiles
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, inside ? last : 0u);
    if (lane == 0 && wmax) atomicMax(tile_nproc + tile, wmax);   // zeroed by the launcher
    __syncthreads();
    if (tid == 0) {
        __threadfence();   // one fence, after the barrier: cumulative over the eight warps' reductions above
        s_last = atomicAdd(&hdr_rw->blend_done, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const uint32_t *np = tile_nproc;
    order_tiles_descending(num_tiles, [np](int t) { return __ldcg(np + t); }, order_bwd, s_bins);
}

// ---- backward -----------------------------------------------------------------------------------------------------------
// All nine (ten) sums a (warp, Gaussian) pair owes the accumulator are LINEAR in two per-pixel scalars: the six raw moments
// are sums over the block's pixels of p = G dL/dalpha times (1, dx, dy, dx^2, dx dy, dy^2) -- weights that depend on the
// pixel position only -- and the colour / inverse-depth derivatives are sums of aT = alpha T times dL/dpixel, which does
// not depend on the Gaussian at all.  So the pixel loop (phase 1) only produces p and aT and parks them in shared memory,
// [Gaussian][pixel]; every NB Gaussians the warp turns around (phase 2): lane (h, g) takes Gaussian g and the 16 pixels of
// half h, forms the ten weighted sums in registers, the two halves meet with one shuffle per sum, and 16 lanes issue
// vector reductions (RED.ADD.v4.F32) into 16 different accumulator rows.  Shared memory is the transposer: ~18 warp
// instructions per Gaussian instead of the ~41 of a register butterfly (31 shuffles + 62 selects + 31 adds per three
// Gaussians) that the kernel used before.
constexpr int NB = 16;             // Gaussians per turn-around
constexpr int NB_PITCH = 33;       // floats per Gaussian row: lanes of both halves hit 32 different banks
struct BwdWarpSmem {
    float4 ring_a[RING];           // x, y, Gaussian id (bits), forward list position (bits)
    float4 ring_co[RING];          // pre-scaled conic, opacity
    float4 ring_feat[RING];        // r, g, b, 1/depth
    float4 meta[NB];               // x, y, Gaussian id (bits) of the parked Gaussians
    float4 dpix[32];               // dL/dpixel (r, g, b, inverse depth) of the block's pixels
    float p[NB * NB_PITCH];
    float aT[NB * NB_PITCH];
};

template <bool INVDEPTH>
__device__ __forceinline__ void backward_turn_around(BwdWarpSmem &sm, int n, uint32_t lane, const WarpBlock &wb,
                                                     float *__restrict__ grad_acc) {
    const int g = (int)(lane & 15u), h = (int)(lane >> 4);
    float S0 = 0.f, Sx = 0.f, Sy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
    const float4 m = sm.meta[g];
    if (g < n) {
        // pixel q = 16 h + l of the block sits at (x0 + l % 8, y0 + 2 h + l / 8);  d = centre - pixel
        const float X0 = m.x - wb.x0, Y0 = m.y - (wb.y0 + (float)(2 * h));
        const float *pp = sm.p + g * NB_PITCH + 16 * h, *pa = sm.aT + g * NB_PITCH + 16 * h;
        const float4 *pw = sm.dpix + 16 * h;
#pragma unroll
        for (int l = 0; l < 16; l++) {
            const float pv = pp[l], av = pa[l];
            const float4 w = pw[l];
            const float dx = X0 - (float)(l & 7), dy = Y0 - (float)(l >> 3);
            const float pdx = pv * dx, pdy = pv * dy;
            S0 += pv; Sx += pdx; Sy += pdy;
            Sxx += pdx * dx; Sxy += pdx * dy; Syy += pdy * dy;
            c0 += av * w.x; c1 += av * w.y; c2 += av * w.z;
            if (INVDEPTH) c3 += av * w.w;
        }
    }
    S0 += __shfl_xor_sync(0xffffffffu, S0, 16); Sx += __shfl_xor_sync(0xffffffffu, Sx, 16);
    Sy += __shfl_xor_sync(0xffffffffu, Sy, 16); Sxx += __shfl_xor_sync(0xffffffffu, Sxx, 16);
    Sxy += __shfl_xor_sync(0xffffffffu, Sxy, 16); Syy += __shfl_xor_sync(0xffffffffu, Syy, 16);
    c0 += __shfl_xor_sync(0xffffffffu, c0, 16); c1 += __shfl_xor_sync(0xffffffffu, c1, 16);
    c2 += __shfl_xor_sync(0xffffffffu, c2, 16);
    if (INVDEPTH) c3 += __shfl_xor_sync(0xffffffffu, c3, 16);
    if (h == 0 && g < n) {
        float *row = grad_acc + (size_t)__float_as_uint(m.z) * GRAD_ACC_FLOATS;
        red_add_v4(row, Sx, Sy, Sxx, Sxy);
        red_add_v4(row + 4, Syy, S0, c0, c1);
        if (INVDEPTH) red_add_v2(row + 8, c2, c3);
        else atomicAdd(row + 8, c2);
    }
}

template <bool INVDEPTH>
__global__ void __launch_bounds__(TILE_PIX, BWD_MIN_CTAS) blend_backward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const float *__restrict__ bg, const float2 *__restrict__ xy,
    const float4 *__restrict__ conic_opacity, const float *__restrict__ colors, const float *__restrict__ depths,
    const float *__restrict__ final_T, const uint32_t *__restrict__ n_contrib, const float *__restrict__ dL_dpix,
    const float *__restrict__ dL_dinvdepth_pix, float *__restrict__ grad_acc, int grid_x,
    const uint32_t *__restrict__ order_bwd) {
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    BwdWarpSmem &sm = reinterpret_cast<BwdWarpSmem *>(s_raw)[warp];

    const uint32_t tile = order_bwd[blockIdx.x];   // deepest contributor first
    const WarpBlock wb = warp_block(tid, (int)(tile % (uint32_t)grid_x), (int)(tile / (uint32_t)grid_x));
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    if (range.y <= range.x) return;
    const uint32_t *point_list = (hdr->final_sel ? vals1 : vals0) + range.x;
    const size_t pid = (size_t)py * W + px, plane = (size_t)H * W;

    const float T_final = inside ? final_T[pid] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pid] : 0u;
    // this warp only walks the prefix of the list that reached one of ITS pixels in the forward, back to front
    const uint32_t n_proc = __reduce_max_sync(0xffffffffu, last_contributor);
    if (n_proc == 0) return;
    const uint32_t nchunks = (n_proc + 31u) / 32u;

    float behind = 0.f;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f, gi = 0.f;
    if (inside) {
        g0 = dL_dpix[pid]; g1 = dL_dpix[plane + pid]; g2 = dL_dpix[2 * plane + pid];
        if (INVDEPTH) gi = dL_dinvdepth_pix[pid];
    }
    sm.dpix[lane] = make_float4(g0, g1, g2, gi);
    const float bg_term = T_final * (bg[0] * g0 + bg[1] * g1 + bg[2] * g2);

    uint32_t p_id = 0, id_next = 0;
    float2 p_xy = make_float2(0.f, 0.f);
    float4 p_co = make_float4(0.f, 0.f, 0.f, 0.f), p_feat = make_float4(0.f, 0.f, 0.f, 0.f);
    auto gather = [&](uint32_t id) {
        p_id = id;
        p_xy = xy[id];
        p_co = conic_opacity[id];
        p_feat = make_float4(colors[3 * (size_t)id], colors[3 * (size_t)id + 1], colors[3 * (size_t)id + 2],
                             INVDEPTH ? depths[id] : 1.f   /* raw depth: dividing here would wait for the load */);
    };
    // lane l of step c looks at forward position q = n_proc - 1 - (32 c + l)
    if (lane < n_proc) gather(point_list[n_proc - 1u - lane]);
    if (32u + lane < n_proc) id_next = point_list[n_proc - 33u - lane];
    uint32_t head = 0, tail = 0;
    int nb = 0;                     // Gaussians parked for the next turn-around
    for (uint32_t c = 0; c < nchunks; c++) {
        const uint32_t back = c * 32u + lane;
        bool hit = false;
        if (back < n_proc) hit = footprint_hit(p_xy, p_co, wb);
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) {
            const uint32_t s = (tail + __popc(m & lanemask_lt())) & (RING - 1);
            sm.ring_a[s] = make_float4(p_xy.x, p_xy.y, __uint_as_float(p_id), __uint_as_float(n_proc - 1u - back));
            sm.ring_co[s] = prescale_conic(p_co);
            sm.ring_feat[s] = make_float4(p_feat.x, p_feat.y, p_feat.z, INVDEPTH ? __fdividef(1.f, p_feat.w) : 0.f);
        }
        tail += __popc(m);
        if (lane == 0) { STAT_ADD(4, 1); STAT_ADD(5, __popc(m)); }
        if (back + 32u < n_proc) gather(id_next);
        if (back + 64u < n_proc) id_next = point_list[n_proc - 65u - back];
        __syncwarp();
        const bool flush = c + 1u == nchunks;
        // phase 1, four ring entries at a time (head stays a multiple of four: a group never wraps): the forward's
        // alphas, evaluated independently, then the transmittance recurrence entry by entry
        while (tail - head >= 4u || (flush && tail != head)) {
            const uint32_t avail = min(4u, tail - head);
            const uint32_t base = head & (RING - 1);
            float G_[4], al[4], ex[4], ey[4], eid[4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 e = sm.ring_a[base + u];
                ex[u] = e.x; ey[u] = e.y; eid[u] = e.z;
                const float dx = e.x - pxf, dy = e.y - pyf;
                const float4 co = sm.ring_co[base + u];   // pre-scaled (log2 domain): exactly the forward's alpha
                const float power = dx * (co.x * dx + co.y * dy) + co.z * dy * dy;
                const float G = fast_exp2(power);
                const float alpha = fminf(0.99f, co.w * G);
                ok[u] = (uint32_t)u < avail && __float_as_uint(e.w) < last_contributor && !(power > 0.f) &&
                        alpha >= 1.f / 255.f;
                G_[u] = G; al[u] = alpha;
            }
            head += avail;
            if (lane == 0) STAT_ADD(6, 1);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                if (!__any_sync(0xffffffffu, ok[u])) continue;   // warp-uniform: no pixel of the block is touched
                const float4 f = sm.ring_feat[base + u];
                const float alpha = al[u];
                const float ra = __fdividef(1.f, 1.f - alpha);
                T = ok[u] ? T * ra : T;                            // transmittance in front of this Gaussian
                const float aT = ok[u] ? alpha * T : 0.f;
                float cg = f.x * g0 + f.y * g1 + f.z * g2;        // (colour . dL/dpixel)
                if (INVDEPTH) cg += f.w * gi;
                // dL/dalpha = (c - accum_rec) . g T - T_final/(1-alpha) (bg . g) with accum_rec T = behind / (1 - alpha),
                // behind = sum over the contributors behind this one of (c . g) alpha T  (the upstream recurrence,
                // contracted with g before instead of after)
                const float dL_dalpha = cg * T - (behind + bg_term) * ra;
                behind += cg * aT;
                sm.p[nb * NB_PITCH + (int)lane] = ok[u] ? G_[u] * dL_dalpha : 0.f;
                sm.aT[nb * NB_PITCH + (int)lane] = aT;
                if (lane == 0) sm.meta[nb] = make_float4(ex[u], ey[u], eid[u], 0.f);
                if (++nb == NB) {
                    __syncwarp();
                    backward_turn_around<INVDEPTH>(sm, NB, lane, wb, grad_acc);
                    __syncwarp();
                    nb = 0;
                }
            }
        }
    }
    if (nb) {
        __syncwarp();
        backward_turn_around<INVDEPTH>(sm, nb, lane, wb, grad_acc);
    }
}

}  // namespace

#ifdef D3GA_BLEND_STATS
extern "C" int d3ga_debug_blend_trace(void *p) {
    unsigned long long *v = (unsigned long long *)p;
    return cudaMemcpyToSymbol(g_blend_trace, &v, sizeof(v)) == cudaSuccess ? 0 : -2;
}
extern "C" int d3ga_debug_blend_stats(unsigned long long *out, int reset) {
    if (cudaMemcpyFromSymbol(out, g_blend_stats, sizeof(g_blend_stats)) != cudaSuccess) return -2;
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (cudaMemcpyToSymbol(g_blend_stats, z, sizeof(z)) != cudaSuccess) return -2;
    }
    return 0;
}
#endif

int launch_blend_forward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                         const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, uint8_t *image,
                         const ImgLayout &IL, float *out_color, float *out_invdepth, cudaStream_t stream) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, num_tiles = gx * gy;
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    RasterHeader *hdr_rw = const_cast<RasterHeader *>(hdr);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const float *depth = reinterpret_cast<const float *>(geom + GL.depth);
    float *final_T = reinterpret_cast<float *>(image + IL.final_T);
    uint32_t *n_contrib = reinterpret_cast<uint32_t *>(image + IL.n_contrib);
    const uint32_t *order_fwd = reinterpret_cast<const uint32_t *>(image + IL.order_fwd);
    uint32_t *order_bwd = reinterpret_cast<uint32_t *>(image + IL.order_bwd);
    uint32_t *tile_nproc = reinterpret_cast<uint32_t *>(image + IL.tile_nproc);
    ProfScope prof(K_BLEND_FWD, stream);
    {
        D3GA_CUDA(cudaMemsetAsync(tile_nproc, 0, sizeof(uint32_t) * (size_t)num_tiles, stream));
        if (out_invdepth)
            blend_forward_kernel<true><<<num_tiles * (8 / FWD_WARPS), FWD_WARPS * 32, 0, stream>>>(
                hdr, ranges, v0, v1, W, H, xy, co, colors, depth, background, final_T, n_contrib, out_color, out_invdepth,
                hdr_rw, gx, num_tiles, order_fwd, tile_nproc, order_bwd);
        else
            blend_forward_kernel<false><<<num_tiles * (8 / FWD_WARPS), FWD_WARPS * 32, 0, stream>>>(
                hdr, ranges, v0, v1, W, H, xy, co, colors, depth, background, final_T, n_contrib, out_color, nullptr,
                hdr_rw, gx, num_tiles, order_fwd, tile_nproc, order_bwd);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

int launch_blend_backward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                          const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, const uint8_t *image,
                          const ImgLayout &IL, const float *dL_dpix, const float *dL_dinvdepth_pix, float *grad_acc,
                          cudaStream_t stream) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, num_tiles = gx * gy;
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const float *depth = reinterpret_cast<const float *>(geom + GL.depth);
    const float *final_T = reinterpret_cast<const float *>(image + IL.final_T);
    const uint32_t *n_contrib = reinterpret_cast<const uint32_t *>(image + IL.n_contrib);
    const uint32_t *order_bwd = reinterpret_cast<const uint32_t *>(image + IL.order_bwd);
    const size_t smem = sizeof(BwdWarpSmem) * (TILE_PIX / 32);
    static std::atomic<uint64_t> attr_devices{0};
    const int arc = once_per_device(attr_devices, [&]() -> int {
        D3GA_CUDA(cudaFuncSetAttribute(blend_backward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        D3GA_CUDA(cudaFuncSetAttribute(blend_backward_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        return D3GA_OK;
    });
    if (arc) return arc;
    ProfScope prof(K_BLEND_BWD, stream);
    {
        if (dL_dinvdepth_pix)
            blend_backward_kernel<true><<<num_tiles, TILE_PIX, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, background, xy, co, colors, depth, final_T, n_contrib, dL_dpix,
                dL_dinvdepth_pix, grad_acc, gx, order_bwd);
        else
            blend_backward_kernel<false><<<num_tiles, TILE_PIX, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, background, xy, co, colors, depth, final_T, n_contrib, dL_dpix, nullptr,
                grad_acc, gx, order_bwd);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
