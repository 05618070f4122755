// raster_blend.cu -- rows R6 (alpha blending forward) and R7 (its backward) of SURVEY.md section 8.
//
// Replaces renderCUDA forward / backward inside diff_gaussian_rasterization._C (reached from renderer.py:130-141
// and from autograd when models/trainer.py:186-192 calls loss.backward()).
//
// One CTA per 16x16 tile (longest list first), one thread per pixel, one warp per 8x4 pixel block.  Each warp walks the
// tile's depth-sorted list ON ITS OWN, 32 entries per step (one per lane), with no CTA barrier in the loop (with batches
// staged by the whole CTA a third of all warp-stall samples sat on that barrier: per-warp culling gives the eight warps
// of a tile very different amounts of work).  A warp whose pixels are all finished simply stops.
//
// What a list entry costs (round 2).  The per-Gaussian gather is ONE 32-byte sector: the SplatRecord the preprocess
// writes (centre, opacity, conic pre-scaled to the log2 domain, 1/depth and the footprint bound lim2), two 16-byte loads,
// fetched a step ahead in registers.  The footprint test of the warp's pixel block (footprint_hit) needs nothing else --
// no logarithm, no determinant per (warp, entry).  Only the SURVIVORS (28 % on the 1 M-Gaussian avatar) are staged: their
// record goes into a per-warp ring in shared memory and their colour is copied global -> shared by cp.async straight
// from the caller's [N,3] tensor (three 4-byte copies; no register round trip, no wait: the ring is consumed one step
// later).  The forward writes its survivor ballots -- one word per (chunk, warp), 1 B per duplicate -- and the BACKWARD
// REPLAYS them: it never tests a footprint and never touches a Gaussian that did not survive; its survivors' records and
// colours all arrive by cp.async.
//   [TMA note: cp.async.bulk needs contiguous sources; a tile's survivors are an index gather of 32-byte records, so the
//    per-thread cp.async (LDGSTS) is the asynchronous copy that fits.  A bulk copy would need the records re-packed in
//    list order first: +64 B per duplicate of streaming writes and reads for lists of which 30 % (forward) and 14 %
//    (backward) are ever visited.]
//
// Forward semantics (same order of tests as upstream): skip power > 0, alpha = min(.99, o * exp(power)), skip
// alpha < 1/255, stop the pixel when T * (1 - alpha) < 1e-4, out = C + T * bg.  alpha is evaluated with one ex2.approx on
// the pre-scaled conic; when it lands within 2e-5 (relative) of the 1/255 cut the CUT -- not the value -- is decided again
// in upstream's operation order with a correctly rounded exponential (exact_cut; taken ~1e-6 of the time), so that the
// set of contributors is the reference's and not a function of this kernel's rounding.  The backward decides identically.
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
#include <math_constants.h>

#include "common.cuh"

namespace d3ga {

namespace {

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_rcp(float x) {   // one MUFU (__fdividef(1, x) adds a denormal rescue: 6 instructions)
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// Asynchronous global -> shared copies (LDGSTS): the data never passes through registers and the issuing lane does not
// wait; cp_async_wait_all() + __syncwarp() make a warp's earlier copies visible to all its lanes.
__device__ __forceinline__ void cp_async_16(void *smem_dst, const void *gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_16_l2(void *smem_dst, const void *gmem_src) {   // not allocated in L1
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src) : "memory");
}
// the backward's survivor copies bypass L1 (measured: 162.0 -> 159.6 us; its L1 hit rate was 25 % with ~30 KB of L1 left)
#ifdef BWD_CP_L1
#define bwd_cp_async_16 cp_async_16
#else
#define bwd_cp_async_16 cp_async_16_l2
#endif
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
#ifdef FWD_BULK_STAGE
// The measured bulk-copy variant of the forward's staging (-DFWD_BULK_STAGE): the copies go
// through the bulk asynchronous copy unit (cp.async.bulk, the 1-D form of TMA) and complete on an mbarrier per stage
// instead of through LDGSTS and commit groups.  Local memory only: source and destination are this GPU's.
__device__ __forceinline__ void bulk_copy(void *smem_dst, const void *gmem_src, uint32_t bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra WAIT_LOOP;\n"
        "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

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
constexpr float CUT = 1.f / 255.f;
constexpr float CUT_BAND = 2e-5f;            // |255 alpha - 1| below which the cut is decided by exact_cut()
#ifndef FWD_MIN_CTAS
#define FWD_MIN_CTAS 3                       // 80 registers, no spills: 165 us; 4 CTAs/SM (64 registers, spills): 181 us
#endif
#ifndef BWD_MIN_CTAS
#define BWD_MIN_CTAS 3
#endif
#ifdef D3GA_BLEND_STATS
// development counters: [0..3] forward (warp-steps, survivors, groups, exact cuts), [4..7] backward
__device__ unsigned long long g_blend_stats[8];
#define STAT_ADD(i, v) do { const unsigned long long _t = (unsigned long long)(v); if (_t) atomicAdd(&g_blend_stats[i], _t); } while (0)
#else
#define STAT_ADD(i, v) do { } while (0)
#endif
#ifdef D3GA_BLEND_TRACE
// development trace of the forward: per (launch slot, warp) start ns, end ns, list entries walked, survivors blended
__device__ unsigned long long g_fwd_trace[65536 * 8 * 4];
// same for the backward: start ns, end ns, tile << 32 | list entries walked, survivors
__device__ unsigned long long g_bwd_trace[65536 * 8 * 4];
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

// Can the Gaussian reach a pixel centre of the warp's block with alpha >= 1/255?  alpha = o 2^q2(d) with
// q2 = A dx^2 + B dx dy + C dy^2 <= 0, so the question is whether max over the block of q2 >= log2(1 / (255 o)).  q2 is
// concave with its maximum at the centre: when the centre lies outside the block the maximum over the block sits on the
// edge(s) facing it, where q2 is a 1-D parabola, maximised in closed form and clamped to the edge.  lim2 (SplatRecord)
// carries the bound with a relative and an absolute margin, so the test is conservative against rounding (a false hit
// only costs time, the per-pixel test decides); +inf = opacity below the cut, -inf = degenerate conic, never culled.
// Rejects the diagonal, elongated splats the bounding box lets through.
__device__ __forceinline__ bool footprint_hit(const float4 g0, const float4 g1, const WarpBlock &wb) {
    const float xl = wb.x0 - g0.x, xh = wb.x1 - g0.x, yl = wb.y0 - g0.y, yh = wb.y1 - g0.y;
    const bool in_x = xl <= 0.f && xh >= 0.f, in_y = yl <= 0.f && yh >= 0.f;
    float qmax = -CUDART_INF_F;
    if (!in_x) {
        const float dx = xl > 0.f ? xl : xh;
        const float dy = fminf(fmaxf((-0.5f * g1.y * dx) * fast_rcp(g1.z), yl), yh);
        qmax = dx * (g1.x * dx + g1.y * dy) + g1.z * dy * dy;
    }
    if (!in_y) {
        const float dy = yl > 0.f ? yl : yh;
        const float dx = fminf(fmaxf((-0.5f * g1.y * dy) * fast_rcp(g1.x), xl), xh);
        qmax = fmaxf(qmax, dx * (g1.x * dx + g1.y * dy) + g1.z * dy * dy);
    }
    return g0.z < CUDART_INF_F && ((in_x && in_y) || !(qmax < g0.z));
}

// The 1/255 cut of one (pixel, Gaussian) pair decided the way the oracle (and upstream) decide it: power in upstream's
// operation order -- individually rounded, no contraction -- and a correctly rounded exponential (double exp rounded to
// float; glibc's expf is correctly rounded in all but ~1e-9 of the cases).  Out of line: reached when the fast alpha is
// within CUT_BAND of the cut, about once per million pairs.
__device__ __noinline__ bool exact_cut(const float2 *__restrict__ xy, const float4 *__restrict__ conic_opacity, uint32_t id,
                                       float pxf, float pyf) {
    const float2 c = xy[id];
    const float4 co = conic_opacity[id];
    const float dx = __fsub_rn(c.x, pxf), dy = __fsub_rn(c.y, pyf);
    const float t = __fadd_rn(__fmul_rn(__fmul_rn(co.x, dx), dx), __fmul_rn(__fmul_rn(co.z, dy), dy));
    const float power = __fsub_rn(__fmul_rn(-0.5f, t), __fmul_rn(__fmul_rn(co.y, dx), dy));
    if (power > 0.f) return false;
    const float alpha = fminf(0.99f, __fmul_rn(co.w, (float)exp((double)power)));
    return alpha >= CUT;
}

// The walk of a tile's list is a per-lane software pipeline in shared memory: lane l of a warp owns entry l of every
// 32-entry chunk and keeps the copies of FWD_DEPTH chunks in flight (cp.async: 2 x 16 B of SplatRecord + 16 B of packed
// colour per entry, into a slot only this lane reads and refills -- no cross-lane hazard, no barrier).  Entry ids run two
// more chunks ahead in registers.  The critical path of a step holds no global load: a heavy tile's warp used to wait
// for its one-step-ahead register prefetch on every chunk (ncu: long-scoreboard the top stall of the forward, the
// busiest SM active for the whole kernel, the average SM for half of it).
#ifndef FWD_DEPTH
#define FWD_DEPTH 2   // 2 and 3 run the kernel equally fast; 2 leaves other views' kernels more shared memory (4: L1 starves)
#endif
// NF = float4s per entry: 3, or 4 for a PAIR render (two images of the same Gaussians in one walk: the second colour)
template <int NF>
struct FwdWarpSmemT {
    float4 ring[RING][NF];                // staged survivors: [0] x, y, list position + 1 (bits), -   [1] conic, opacity
                                          //                   [2] r, g, b, 1/depth   ([3] second colour)
#ifndef FWD_BULK_STAGE
    float4 stage[FWD_DEPTH][NF][32];      // chunks in flight: [chunk][g0 | g1 | colour (| second colour)][lane]
#else
    SplatRecord rec[FWD_DEPTH][32];       // chunks in flight: a lane's record contiguous (one 32-byte bulk copy)
    float4 col[FWD_DEPTH][32];
    unsigned long long bar[FWD_DEPTH];    // one mbarrier per stage: completes when the chunk's bytes have landed
#endif
};

// PAIR: a second image of the SAME Gaussians with other colours and another background (the silhouette pass of a D3GA
// training frame, models/trainer.py:102-110) is blended in the same walk: same alphas, same transmittance, same
// contributor set -- one more 16-byte colour per survivor and three more accumulators per pixel.
#ifndef FWD_PAIR_MIN_CTAS
#define FWD_PAIR_MIN_CTAS 2   // 99 registers, no spills: 180 us at 1 M / 1024^2; 3 CTAs (80 registers, 52 B of spills, 192 KB of shared memory): 292 us
#endif
template <bool INVDEPTH, bool PAIR>
__global__ void __launch_bounds__(FWD_WARPS * 32, PAIR ? FWD_PAIR_MIN_CTAS : FWD_MIN_CTAS) blend_forward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const SplatRecord *__restrict__ splat,
    const float4 *__restrict__ feat, const float2 *__restrict__ xy, const float4 *__restrict__ conic_opacity,
    const float *__restrict__ bg, float *__restrict__ final_T, uint32_t *__restrict__ n_contrib,
    float *__restrict__ out_color, float *__restrict__ out_invdepth, RasterHeader *__restrict__ hdr_rw, int grid_x,
    int num_tiles, const uint32_t *__restrict__ order_fwd, uint32_t *__restrict__ tile_nproc,
    uint32_t *__restrict__ order_bwd, uint32_t *__restrict__ ballots, const float4 *__restrict__ feat_b,
    const float *__restrict__ bg_b, float *__restrict__ out_color_b) {
    constexpr int NF = PAIR ? 4 : 3;
    using FwdWarpSmem = FwdWarpSmemT<NF>;
    extern __shared__ __align__(16) unsigned char s_raw[];
    __shared__ uint32_t s_bins[257];
    __shared__ uint32_t s_last;
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    FwdWarpSmem &sm = reinterpret_cast<FwdWarpSmem *>(s_raw)[warp];

    // The blend loop below is branch-free: entries of a partly filled last group are multiplied by a weight of zero
    // rather than skipped, so a ring slot that was never written must not hold a NaN or Inf bit pattern.
    for (int i = (int)lane; i < RING * NF; i += 32) (&sm.ring[0][0])[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    const uint32_t tile = order_fwd[blockIdx.x];   // longest Gaussian list first
    WarpBlock wb = warp_block((int)warp, (int)lane, (int)(tile % (uint32_t)grid_x), (int)(tile / (uint32_t)grid_x));
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const float blk_x0 = wb.x0, blk_y0 = wb.y0;
    // The footprint test runs against the bounding box of the block's pixels that are STILL BEING BLENDED, not against
    // the whole 8x4 block: on a silhouette tile the pixels inside the body saturate early and the rest of the (long) list
    // only matters to the few pixels outside it, so most of what follows no longer survives -- in this kernel and, through
    // the ballots, in the backward.  An entry culled this way contributes nothing: a finished pixel takes no entry, an
    // active pixel lies inside the box the (conservative) test was made against.  active_box() is called when the set of
    // active pixels has changed (bit l of the mask = lane l = pixel (l % 8, l / 8) of the block).
    uint32_t act_mask = 0xffffffffu;
    auto active_box = [&](uint32_t m) {
        act_mask = m;
        if (m == 0u) return;
        const uint32_t cols = (m | (m >> 8) | (m >> 16) | (m >> 24)) & 0xffu;        // columns with an active pixel
        const int xlo = __ffs((int)cols) - 1, xhi = 31 - __clz((int)cols);
        const int ylo = (__ffs((int)m) - 1) >> 3, yhi = (31 - __clz((int)m)) >> 3;
        wb.x0 = blk_x0 + (float)xlo; wb.x1 = blk_x0 + (float)xhi;
        wb.y0 = blk_y0 + (float)ylo; wb.y1 = blk_y0 + (float)yhi;
    };
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    const uint32_t *point_list = (hdr->final_sel ? vals1 : vals0) + range.x;
    const uint32_t count = range.y - range.x;
    const uint32_t nchunks = (count + 31u) / 32u;
    uint32_t *ballot_w = ballots + ((size_t)(range.x >> 5) + tile) * 8 + warp;   // + 8 per chunk

    bool done = !inside;
    float T = 1.f, C0 = 0.f, C1 = 0.f, C2 = 0.f, invd = 0.f;
    float D0 = 0.f, D1 = 0.f, D2 = 0.f;     // PAIR: the second image
    uint32_t last = 0;
    uint32_t head = 0, tail = 0;
    active_box(__ballot_sync(0xffffffffu, !done));
    bool all_done = act_mask == 0u;
#ifdef D3GA_BLEND_TRACE
    const unsigned long long tr_t0 = global_ns();
    uint32_t tr_chunks = 0;
#endif

    // this lane's copies of one chunk entry: record + colour -> its private slot of stage `st`
#ifndef FWD_BULK_STAGE
    auto fetch = [&](uint32_t st, uint32_t id) {
        cp_async_16(&sm.stage[st][0][lane], &splat[id].g0);
        cp_async_16(&sm.stage[st][1][lane], &splat[id].g1);
        cp_async_16(&sm.stage[st][2][lane], feat + id);
        if (PAIR) cp_async_16(&sm.stage[st][NF - 1][lane], feat_b + id);
    };
#else
    static_assert(!PAIR, "the bulk-copy staging variant has no pair render");
    uint32_t bulk_parity = 0, bulk_pending = 0;   // per stage: the phase to wait for next / a chunk is in flight
    if (lane == 0) {
#pragma unroll
        for (int j = 0; j < FWD_DEPTH; j++) mbar_init(&sm.bar[j], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    fence_proxy_async();
    __syncwarp();
    // chunk `chunk` into stage st: lane 0 arms the barrier with the chunk's bytes, every lane with an entry issues its two
    // bulk copies (the phase cannot complete before lane 0's arrive, whatever order the copies land in)
    auto fetch_chunk = [&](uint32_t st, uint32_t chunk, uint32_t id) {
        const uint32_t first = chunk * 32u;
        if (first >= count) return;
        const uint32_t nvalid = min(32u, count - first);
        fence_proxy_async();                       // the stage was last read through the generic proxy
        if (lane == 0) mbar_expect_tx(&sm.bar[st], 48u * nvalid);
        if (lane < nvalid) {
            bulk_copy(&sm.rec[st][lane], splat + id, 32u, &sm.bar[st]);
            bulk_copy(&sm.col[st][lane], feat + id, 16u, &sm.bar[st]);
        }
        bulk_pending |= 1u << st;
    };
#endif
    // prologue: FWD_DEPTH chunks in flight (one commit group per chunk, empty ones included).  idq[s] is the id of this
    // lane's entry in the next chunk that will be fetched into stage s: it is loaded FWD_DEPTH steps before it is used,
    // into a register that nothing touches in between (the chunk loop is unrolled over the stages, so the queue needs no
    // moves -- a move of the newest id made every step wait for the load of the step before, at DRAM latency in the tail).
    uint32_t idq[FWD_DEPTH];
#pragma unroll
    for (int j = 0; j < FWD_DEPTH; j++) idq[j] = 0;
    if (!all_done) {
#pragma unroll
        for (uint32_t j = 0; j < (uint32_t)FWD_DEPTH; j++) {
#ifndef FWD_BULK_STAGE
            if (j * 32u + lane < count) fetch(j, point_list[j * 32u + lane]);
            cp_async_commit();
#else
            fetch_chunk(j, j, j * 32u + lane < count ? point_list[j * 32u + lane] : 0u);
#endif
        }
#pragma unroll
        for (uint32_t j = 0; j < (uint32_t)FWD_DEPTH; j++)
            if (((uint32_t)FWD_DEPTH + j) * 32u + lane < count) idq[j] = point_list[((uint32_t)FWD_DEPTH + j) * 32u + lane];
    }

    // Blends the staged survivors, four ring entries (a GROUP) at a time; head stays a multiple of four, so a group never
    // wraps.  What bounds the kernel is the serial chain of the busiest warp of the densest tile (ncu: the busiest SM is
    // active for the whole kernel, the average SM for half of it), so the loop is written for LATENCY:
    //  - software pipeline: the alphas of group g + 1 (shared-memory loads, the quadratic form, ex2) are computed in the
    //    same basic block as the recurrence of group g, so the in-order issue interleaves the two chains;
    //  - the transmittance chain is four dependent multiplies p_k = p_{k-1} (1 - a_k); the stop T (1 - a) < 1e-4 is found
    //    beside the chain from the four products (T only ever decreases) and applied as masks.  Same products, same
    //    roundings, same decisions as the entry-by-entry form (and as upstream: subtract, then multiply).
    float ga[4] = {0.f, 0.f, 0.f, 0.f};      // the pending group's alphas (0 = skipped entry) ...
    uint32_t gpos[4] = {0u, 0u, 0u, 0u};     // ... and list positions + 1
    uint32_t gavail = 0;                     // entries of the pending group (0: none pending)
    float gnear = 1.f;
    auto alphas_of = [&](uint32_t h, uint32_t avail, float (&a)[4], uint32_t (&posv)[4], float &nearest) {
        const float4(*g)[NF] = sm.ring + (h & (RING - 1));
        nearest = 1.f;
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const float4 e = g[u][0];
            const float dx = e.x - pxf, dy = e.y - pyf;
            const float4 co = g[u][1];   // pre-scaled: log2 domain
            const float power = dx * (co.x * dx + co.y * dy) + co.z * dy * dy;
            const float alpha = fminf(0.99f, co.w * fast_exp2(power));
            const bool keep = (uint32_t)u < avail && !(power > 0.f) && alpha >= CUT;
            a[u] = keep ? alpha : 0.f;
            nearest = (uint32_t)u < avail ? fminf(nearest, fabsf(fmaf(alpha, 255.f, -1.f))) : nearest;
            posv[u] = __float_as_uint(e.z);
        }
    };
    // ~1e-6 of the pairs: the contributor set must not depend on this kernel's rounding
    auto resolve_cuts = [&](uint32_t h, uint32_t avail, float (&a)[4], const uint32_t (&posv)[4]) {
        const float4(*g)[NF] = sm.ring + (h & (RING - 1));
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const float4 e = g[u][0];
            const float4 co = g[u][1];
            const float dx = e.x - pxf, dy = e.y - pyf;
            const float alpha = fminf(0.99f, co.w * fast_exp2(dx * (co.x * dx + co.y * dy) + co.z * dy * dy));
            if ((uint32_t)u < avail && fabsf(fmaf(alpha, 255.f, -1.f)) < CUT_BAND) {
                a[u] = exact_cut(xy, conic_opacity, point_list[posv[u] - 1u], pxf, pyf) ? alpha : 0.f;
                STAT_ADD(3, 1);
            }
        }
    };
    auto consume = [&](bool flush) {
        auto ready = [&]() { return tail - head >= 4u || (flush && tail != head); };
        if (!ready()) return;
        gavail = min(4u, tail - head);
        alphas_of(head, gavail, ga, gpos, gnear);
        while (true) {
            const uint32_t h = head;
            if (gnear < CUT_BAND) resolve_cuts(h, gavail, ga, gpos);
            head += gavail;
            // ---- one basic block: alphas of the next group (if any) beside the recurrence of this one ----
            const uint32_t navail = ready() ? min(4u, tail - head) : 0u;
            float na[4] = {0.f, 0.f, 0.f, 0.f};
            uint32_t npos[4] = {0u, 0u, 0u, 0u};
            float nnear = 1.f;
            if (navail) alphas_of(head, navail, na, npos, nnear);   // (warp-uniform; a third of the calls end here)
            const float4(*g)[NF] = sm.ring + (h & (RING - 1));
            const float p0 = T;
            const float p1 = p0 * (1.f - ga[0]), p2 = p1 * (1.f - ga[1]), p3 = p2 * (1.f - ga[2]), p4 = p3 * (1.f - ga[3]);
            const bool s0 = done | (p1 < 0.0001f), s1 = s0 | (p2 < 0.0001f), s2 = s1 | (p3 < 0.0001f), s3 = s2 | (p4 < 0.0001f);
            const float w0 = s0 ? 0.f : ga[0] * p0, w1 = s1 ? 0.f : ga[1] * p1, w2 = s2 ? 0.f : ga[2] * p2,
                        w3 = s3 ? 0.f : ga[3] * p3;
            T = s0 ? p0 : (s1 ? p1 : (s2 ? p2 : (s3 ? p3 : p4)));
            done = s3;
            const float4 f0 = g[0][2], f1 = g[1][2], f2 = g[2][2], f3 = g[3][2];
            C0 += f0.x * w0 + f1.x * w1 + f2.x * w2 + f3.x * w3;
            C1 += f0.y * w0 + f1.y * w1 + f2.y * w2 + f3.y * w3;
            C2 += f0.z * w0 + f1.z * w1 + f2.z * w2 + f3.z * w3;
            if (INVDEPTH) invd += f0.w * w0 + f1.w * w1 + f2.w * w2 + f3.w * w3;
            if (PAIR) {
                const float4 e0 = g[0][NF - 1], e1 = g[1][NF - 1], e2 = g[2][NF - 1], e3 = g[3][NF - 1];
                D0 += e0.x * w0 + e1.x * w1 + e2.x * w2 + e3.x * w3;
                D1 += e0.y * w0 + e1.y * w1 + e2.y * w2 + e3.y * w3;
                D2 += e0.z * w0 + e1.z * w1 + e2.z * w2 + e3.z * w3;
            }
            last = w0 != 0.f ? gpos[0] : last;
            last = w1 != 0.f ? gpos[1] : last;
            last = w2 != 0.f ? gpos[2] : last;
            last = w3 != 0.f ? gpos[3] : last;
            {
                const uint32_t still = __ballot_sync(0xffffffffu, !done);
                if (still == 0u) { all_done = true; return; }
                if (still != act_mask) active_box(still);
            }
            if (navail == 0u) return;
#pragma unroll
            for (int u = 0; u < 4; u++) { ga[u] = na[u]; gpos[u] = npos[u]; }
            gavail = navail;
            gnear = nnear;
        }
    };

    for (uint32_t c0 = 0; c0 < nchunks && !all_done; c0 += (uint32_t)FWD_DEPTH) {
#pragma unroll
        for (int st = 0; st < FWD_DEPTH; st++) {
            const uint32_t c = c0 + (uint32_t)st;
            if (c >= nchunks || all_done) break;
            // this lane's entry of chunk c has landed when at most FWD_DEPTH - 1 younger groups are still in flight
#ifndef FWD_BULK_STAGE
            asm volatile("cp.async.wait_group %0;" ::"n"(FWD_DEPTH - 1) : "memory");
#else
            mbar_wait(&sm.bar[st], (bulk_parity >> st) & 1u);
            bulk_parity ^= 1u << st;
            bulk_pending &= ~(1u << st);
#endif
            const uint32_t pos = c * 32u + lane;
            float4 g0 = make_float4(0.f, 0.f, CUDART_INF_F, 0.f), g1 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (pos < count) {
#ifndef FWD_BULK_STAGE
                g0 = sm.stage[st][0][lane];
                g1 = sm.stage[st][1][lane];
#else
                g0 = sm.rec[st][lane].g0;
                g1 = sm.rec[st][lane].g1;
#endif
            }
            const bool hit = pos < count && footprint_hit(g0, g1, wb);
            const uint32_t m = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) {
                ballot_w[(size_t)c * 8] = m;     // replayed by the backward
                STAT_ADD(0, 1); STAT_ADD(1, __popc(m));
            }
            if (hit) {
#ifndef FWD_BULK_STAGE
                const float4 col = sm.stage[st][2][lane];
#else
                const float4 col = sm.col[st][lane];
#endif
                float4 *slot = sm.ring[(tail + __popc(m & lanemask_lt())) & (RING - 1)];
                slot[0] = make_float4(g0.x, g0.y, __uint_as_float(pos + 1u), 0.f);
                slot[1] = make_float4(g1.x, g1.y, g1.z, g0.w);
                slot[2] = make_float4(col.x, col.y, col.z, g1.w);
#ifndef FWD_BULK_STAGE
                if (PAIR) slot[NF - 1] = sm.stage[st][NF - 1][lane];
#endif
            }
            tail += __popc(m);
            // refill the stage with this lane's entry of chunk c + FWD_DEPTH; its id for the chunk after that
            const uint32_t ahead = pos + (uint32_t)FWD_DEPTH * 32u;
#ifndef FWD_BULK_STAGE
            if (ahead < count) fetch((uint32_t)st, idq[st]);
            cp_async_commit();
#else
            __syncwarp();                          // every lane has taken its entry out of the stage
            fetch_chunk((uint32_t)st, c + (uint32_t)FWD_DEPTH, idq[st]);
#endif
            if (ahead + (uint32_t)FWD_DEPTH * 32u < count) idq[st] = point_list[ahead + (uint32_t)FWD_DEPTH * 32u];
            // the list itself is streamed from HBM once per tile: pull its lines into L2 well ahead of the id loads
            if (lane == 0 && pos + 32u * 24u < count)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(point_list + pos + 32u * 24u));
            __syncwarp();
            consume(false);
#ifdef D3GA_BLEND_TRACE
            tr_chunks++;
#endif
        }
    }
    if (!all_done) consume(true);
#ifdef D3GA_BLEND_TRACE
    if (lane == 0 && blockIdx.x < 65536) {
        unsigned long long *o = g_fwd_trace + 4 * ((size_t)blockIdx.x * 8 + warp);
        o[0] = tr_t0; o[1] = global_ns(); o[2] = ((unsigned long long)tile << 32) | tr_chunks; o[3] = ((unsigned long long)count << 32) | head;
    }
#endif
    cp_async_wait_all();   // nothing may still be landing in shared memory when the CTA retires
#ifdef FWD_BULK_STAGE
#pragma unroll
    for (int j = 0; j < FWD_DEPTH; j++)
        if ((bulk_pending >> j) & 1u) mbar_wait(&sm.bar[j], (bulk_parity >> j) & 1u);
#endif
    if (inside) {
        const size_t pid = (size_t)py * W + px, plane = (size_t)H * W;
        final_T[pid] = T;
        n_contrib[pid] = last;
        out_color[pid] = C0 + T * bg[0];
        out_color[plane + pid] = C1 + T * bg[1];
        out_color[2 * plane + pid] = C2 + T * bg[2];
        if (INVDEPTH) out_invdepth[pid] = invd;
        if (PAIR) {
            out_color_b[pid] = D0 + T * bg_b[0];
            out_color_b[plane + pid] = D1 + T * bg_b[1];
            out_color_b[2 * plane + pid] = D2 + T * bg_b[2];
        }
    }
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
template <int NF>
struct BwdWarpSmemT {
    // per staged survivor 48 contiguous bytes, all landing by cp.async:
    //   [0] SplatRecord.g0 as it lies in global memory: x, y, (lim2), opacity   [1] SplatRecord.g1: pre-scaled conic, 1/depth
    //   [2] r, g, b (from the colour tensor), unused   ([3] PAIR render: the second image's colour)
    float4 ring[RING][NF];
    uint2 ring_m[RING];            // Gaussian id, forward list position
    float4 meta[NB];               // x, y, Gaussian id (bits) of the parked Gaussians
    float4 dpix[32];               // dL/dpixel (r, g, b, inverse depth) of the block's pixels
    float p[NB * NB_PITCH];
    float aT[NB * NB_PITCH];
};

// (Out of line: inlined at every place a group can fill the park, the kernel was ~95 KB of code and a tenth of its
// issue slots were lost waiting for instructions.)
template <bool INVDEPTH, typename BwdWarpSmem>
__device__ __noinline__ void backward_turn_around(BwdWarpSmem &sm, int n, uint32_t lane, float bx0, float by0,
                                                  float *__restrict__ grad_acc) {
    const int g = (int)(lane & 15u), h = (int)(lane >> 4);
    float S0 = 0.f, Sx = 0.f, Sy = 0.f, Sxx = 0.f, Sxy = 0.f, Syy = 0.f, c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
    const float4 m = sm.meta[g];
    const uint32_t gid = __float_as_uint(m.z);
    if (g < n) {
        // pixel q = 16 h + l of the block sits at (x0 + l % 8, y0 + 2 h + l / 8).  The moments are first taken about the
        // corner of the half-block with the INTEGER pixel offsets (lx, ly) as weights -- compile-time constants of the
        // unrolled loop, ly in {0, 1}: ~3.7 instead of 10 instructions per pixel -- and then moved to the Gaussian's
        // centre, d = (X0 - lx, Y0 - ly):  sum p dx = X0 M0 - Mx,  sum p dx^2 = X0^2 M0 - 2 X0 Mx + Mxx, ...
        const float *pp = sm.p + g * NB_PITCH + 16 * h, *pa = sm.aT + g * NB_PITCH + 16 * h;
        const float4 *pw = sm.dpix + 16 * h;
        float M0 = 0.f, Mx = 0.f, My = 0.f, Mxx = 0.f, Mxy = 0.f;
#pragma unroll
        for (int l = 0; l < 16; l++) {
            const float pv = pp[l], av = pa[l];
            const float4 w = pw[l];
            const float lx = (float)(l & 7);
            M0 += pv;
            if ((l & 7) != 0) { Mx += pv * lx; Mxx += pv * (lx * lx); }
            if ((l >> 3) != 0) {
                My += pv;                                    // ly = 1 (and ly^2 = ly: Myy = My)
                if ((l & 7) != 0) Mxy += pv * lx;
            }
            c0 += av * w.x; c1 += av * w.y; c2 += av * w.z;
            if (INVDEPTH) c3 += av * w.w;
        }
        const float X0 = m.x - bx0, Y0 = m.y - (by0 + (float)(2 * h));
        S0 = M0;
        Sx = X0 * M0 - Mx;
        Sy = Y0 * M0 - My;
        Sxx = X0 * Sx - (X0 * Mx - Mxx);                     // X0^2 M0 - 2 X0 Mx + Mxx
        Sxy = Y0 * Sx - (X0 * My - Mxy);                     // X0 Y0 M0 - Y0 Mx - X0 My + Mxy
        Syy = Y0 * Sy - (Y0 * My - My);                      // Y0^2 M0 - 2 Y0 My + Myy
    }
    S0 += __shfl_xor_sync(0xffffffffu, S0, 16); Sx += __shfl_xor_sync(0xffffffffu, Sx, 16);
    Sy += __shfl_xor_sync(0xffffffffu, Sy, 16); Sxx += __shfl_xor_sync(0xffffffffu, Sxx, 16);
    Sxy += __shfl_xor_sync(0xffffffffu, Sxy, 16); Syy += __shfl_xor_sync(0xffffffffu, Syy, 16);
    c0 += __shfl_xor_sync(0xffffffffu, c0, 16); c1 += __shfl_xor_sync(0xffffffffu, c1, 16);
    c2 += __shfl_xor_sync(0xffffffffu, c2, 16);
    if (INVDEPTH) c3 += __shfl_xor_sync(0xffffffffu, c3, 16);
    if (h == 0 && g < n) {
        float *row = grad_acc + (size_t)gid * GRAD_ACC_FLOATS;
        red_add_v4(row, Sx, Sy, Sxx, Sxy);
        red_add_v4(row + 4, Syy, S0, c0, c1);
        if (INVDEPTH) red_add_v2(row + 8, c2, c3);
        else atomicAdd(row + 8, c2);
    }
}

// PAIR: the backward of a pair render (see blend_forward_kernel): the second image's dL/dpixel joins the first's in
// dL/dalpha -- (c . g) becomes (c . g) + (c_b . g_b), the background term likewise -- so geometry and opacity receive the
// gradient of both images from ONE walk; the second image's colours (constants in D3GA: cage_net.py:243) get no gradient.
template <bool INVDEPTH, bool PAIR>
__global__ void __launch_bounds__(TILE_PIX, BWD_MIN_CTAS) blend_backward_kernel(
    const RasterHeader *__restrict__ hdr, const uint2 *__restrict__ ranges, const uint32_t *__restrict__ vals0,
    const uint32_t *__restrict__ vals1, int W, int H, const float *__restrict__ bg,
    const SplatRecord *__restrict__ splat, const float4 *__restrict__ feat, const float2 *__restrict__ xy,
    const float4 *__restrict__ conic_opacity, const float *__restrict__ final_T, const uint32_t *__restrict__ n_contrib,
    const float *__restrict__ dL_dpix, const float *__restrict__ dL_dinvdepth_pix, float *__restrict__ grad_acc,
    int grid_x, const uint32_t *__restrict__ order_bwd, const uint32_t *__restrict__ ballots,
    const float4 *__restrict__ feat_b, const float *__restrict__ bg_b, const float *__restrict__ dL_dpix_b) {
    constexpr int NF = PAIR ? 4 : 3;
    using BwdWarpSmem = BwdWarpSmemT<NF>;
    extern __shared__ __align__(16) unsigned char s_raw[];
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    BwdWarpSmem &sm = reinterpret_cast<BwdWarpSmem *>(s_raw)[warp];

#ifdef D3GA_BLEND_TRACE
    const unsigned long long tr_t0 = global_ns();
    uint32_t tr_surv = 0;
#endif
    const uint32_t tile = order_bwd[blockIdx.x];   // deepest contributor first
    const WarpBlock wb = warp_block(tid, (int)(tile % (uint32_t)grid_x), (int)(tile / (uint32_t)grid_x));
    const int px = wb.px, py = wb.py;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    uint2 range = make_uint2(0, 0);
    if (!hdr->overflow) range = ranges[tile];
    if (range.y <= range.x) return;
    const uint32_t *point_list = (hdr->final_sel ? vals1 : vals0) + range.x;
    const uint32_t *ballot_w = ballots + ((size_t)(range.x >> 5) + tile) * 8 + warp;   // + 8 per forward chunk
    const size_t pid = (size_t)py * W + px, plane = (size_t)H * W;

    const float T_final = inside ? final_T[pid] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pid] : 0u;
    // this warp only walks the prefix of the list that reached one of ITS pixels in the forward, back to front
    const uint32_t n_proc = __reduce_max_sync(0xffffffffu, last_contributor);
    if (n_proc == 0) return;

    float behind = 0.f;
    float g0 = 0.f, g1 = 0.f, g2 = 0.f, gi = 0.f;
    if (inside) {
        g0 = dL_dpix[pid]; g1 = dL_dpix[plane + pid]; g2 = dL_dpix[2 * plane + pid];
        if (INVDEPTH) gi = dL_dinvdepth_pix[pid];
    }
    sm.dpix[lane] = make_float4(g0, g1, g2, gi);
    float h0 = 0.f, h1 = 0.f, h2 = 0.f;     // PAIR: dL/dpixel of the second image
    if (PAIR && inside) { h0 = dL_dpix_b[pid]; h1 = dL_dpix_b[plane + pid]; h2 = dL_dpix_b[2 * plane + pid]; }
    float bg_term = T_final * (bg[0] * g0 + bg[1] * g1 + bg[2] * g2);
    if (PAIR) bg_term += T_final * (bg_b[0] * h0 + bg_b[1] * h1 + bg_b[2] * h2);

    // The forward's chunks, last one first; lane l of a step looks at forward position q = 32 fc + 31 - l, so that the
    // compaction (lower lanes first) keeps the back-to-front order.  The forward's ballot of (chunk, warp) says which of
    // the 32 entries reached this warp's pixel block: nothing else is gathered, no footprint is tested again.
    const int top = (int)((n_proc - 1u) >> 5);
    const uint32_t top_mask = ((n_proc - 1u) & 31u) == 31u ? 0xffffffffu : ((2u << ((n_proc - 1u) & 31u)) - 1u);
    // ballots and ids run two steps ahead in registers
    uint32_t m_next = __ldg(ballot_w + (size_t)top * 8) & top_mask;
    uint32_t id_next = point_list[min(32u * (uint32_t)top + 31u - lane, n_proc - 1u)];
    uint32_t m_next2 = 0, id_next2 = 0;
    if (top > 0) {
        m_next2 = __ldg(ballot_w + (size_t)(top - 1) * 8);
        id_next2 = point_list[32u * (uint32_t)(top - 1) + 31u - lane];
    }
    uint32_t head = 0, tail = 0;
    int nb = 0;                     // Gaussians parked for the next turn-around

    // phase 1, four ring entries at a time (head stays a multiple of four: a group never wraps): the forward's
    // alphas, evaluated independently, then the transmittance recurrence entry by entry
    auto consume = [&](bool flush) {
        while (tail - head >= 4u || (flush && tail != head)) {
            const uint32_t avail = min(4u, tail - head);
            const uint32_t base = head & (RING - 1);
            const float4(*g)[NF] = sm.ring + base;
            // al = the forward's alpha where this pixel took the Gaussian, else 0 (alpha >= 1/255 there, so "taken" is
            // al != 0); Gk = G there, else 0.  A pair that is not taken then drops out of every sum below by itself:
            // ra = 1 leaves T alone, alpha T = 0, G dL/dalpha = 0 -- one select in the recurrence instead of four.
            float Gk[4], al[4];
            uint32_t near = 0;
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const float4 e = g[u][0];     // x, y, -, opacity
                const float4 co = g[u][1];    // pre-scaled conic (log2 domain): exactly the forward's alpha
                const uint32_t q = sm.ring_m[base + u].y;
                const float dx = e.x - pxf, dy = e.y - pyf;
                const float power = dx * (co.x * dx + co.y * dy) + co.z * dy * dy;
                const float G = fast_exp2(power);
                const float alpha = fminf(0.99f, e.w * G);
                const bool mine = (uint32_t)u < avail && q < last_contributor;
                const bool ok = mine && !(power > 0.f) && alpha >= CUT;
                near |= (mine && fabsf(fmaf(alpha, 255.f, -1.f)) < CUT_BAND) ? (1u << u) : 0u;
                Gk[u] = ok ? G : 0.f;
                al[u] = ok ? alpha : 0.f;
            }
            if (near) {   // the same decision as the forward's (see exact_cut)
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (!((near >> u) & 1u)) continue;
                    const float4 e = g[u][0];
                    const float4 co = g[u][1];
                    const float dx = e.x - pxf, dy = e.y - pyf;
                    const float G = fast_exp2(dx * (co.x * dx + co.y * dy) + co.z * dy * dy);
                    const bool ok = exact_cut(xy, conic_opacity, sm.ring_m[base + u].x, pxf, pyf);
                    Gk[u] = ok ? G : 0.f;
                    al[u] = ok ? fminf(0.99f, e.w * G) : 0.f;
                }
            }
            head += avail;
            if (lane == 0) STAT_ADD(6, 1);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool taken = al[u] != 0.f;
                if (!__any_sync(0xffffffffu, taken)) continue;     // warp-uniform: no pixel of the block is touched
                const float4 f = g[u][2];
                const float alpha = al[u];
                const float ra = taken ? fast_rcp(1.f - alpha) : 1.f;   // 1 - alpha >= 0.01
                T *= ra;                                           // transmittance in front of this Gaussian
                const float aT = alpha * T;
                float cg = f.x * g0 + f.y * g1 + f.z * g2;        // (colour . dL/dpixel)
                if (INVDEPTH) cg += g[u][1].w * gi;
                if (PAIR) { const float4 fb = g[u][NF - 1]; cg += fb.x * h0 + fb.y * h1 + fb.z * h2; }
                // dL/dalpha = (c - accum_rec) . g T - T_final/(1-alpha) (bg . g) with accum_rec T = behind / (1 - alpha),
                // behind = sum over the contributors behind this one of (c . g) alpha T  (the upstream recurrence,
                // contracted with g before instead of after)
                const float dL_dalpha = cg * T - (behind + bg_term) * ra;
                behind += cg * aT;
                sm.p[nb * NB_PITCH + (int)lane] = Gk[u] * dL_dalpha;
                sm.aT[nb * NB_PITCH + (int)lane] = aT;
                // centre and id for the turn-around, copied out of the ring slot (the slot may be refilled before the
                // turn-around comes: parked entries are the TAKEN ones, any number of skipped ones can lie between them)
                if (lane == 0) {
                    const float4 e = g[u][0];
                    sm.meta[nb] = make_float4(e.x, e.y, __uint_as_float(sm.ring_m[base + u].x), 0.f);
                }
                nb++;
            }
            if (nb > NB - 4) {   // the next group may not fit: one call site per group, not one per entry
                __syncwarp();
                backward_turn_around<INVDEPTH>(sm, nb, lane, wb.x0, wb.y0, grad_acc);
                __syncwarp();
                nb = 0;
            }
        }
    };

    for (int fc = top; fc >= 0; fc--) {
        const uint32_t mf = m_next, id = id_next;
        // a chunk none of whose entries reached this warp's pixels, with less than a group waiting: nothing to blend and
        // nothing to stage -- only the look-ahead moves on (deep in a silhouette tile's list most chunks are like that)
        const bool idle = mf == 0u && tail - head < 4u;
        if (!idle) {
            // the survivors staged by the previous steps (their records have had a whole step to land)
            cp_async_wait_all();
            __syncwarp();
            consume(false);
        }
        m_next = m_next2;
        id_next = id_next2;
        if (fc > 1) {
            m_next2 = __ldg(ballot_w + (size_t)(fc - 2) * 8);
            id_next2 = point_list[32u * (uint32_t)(fc - 2) + 31u - lane];
        }
        const uint32_t m = __brev(mf);                       // bit l = lane l = forward position 32 fc + 31 - l
        if ((m >> lane) & 1u) {
            const uint32_t s = (tail + __popc(m & lanemask_lt())) & (RING - 1);
            bwd_cp_async_16(&sm.ring[s][0], &splat[id].g0);
            bwd_cp_async_16(&sm.ring[s][1], &splat[id].g1);
            bwd_cp_async_16(&sm.ring[s][2], feat + id);
            if (PAIR) bwd_cp_async_16(&sm.ring[s][NF - 1], feat_b + id);
            sm.ring_m[s] = make_uint2(id, 32u * (uint32_t)fc + 31u - lane);
        }
        if (mf != 0u) cp_async_commit();
        tail += __popc(m);
#ifdef D3GA_BLEND_TRACE
        tr_surv += __popc(m);
#endif
        if (lane == 0) { STAT_ADD(4, 1); STAT_ADD(5, __popc(m)); }
    }
    cp_async_wait_all();
    __syncwarp();
    consume(true);
    if (nb) {
        __syncwarp();
        backward_turn_around<INVDEPTH>(sm, nb, lane, wb.x0, wb.y0, grad_acc);
    }
#ifdef D3GA_BLEND_TRACE
    if (lane == 0 && blockIdx.x < 65536) {
        unsigned long long *o = g_bwd_trace + 4 * ((size_t)blockIdx.x * 8 + warp);
        o[0] = tr_t0; o[1] = global_ns(); o[2] = ((unsigned long long)tile << 32) | n_proc; o[3] = tr_surv;
    }
#endif
}

// This render's colours, [N,3] as the API hands them (or the SH colours of the preprocess), as one aligned float4 per
// Gaussian: a list entry's colour is then ONE 16-byte asynchronous copy instead of three 4-byte ones.  28 B per Gaussian.
__global__ void __launch_bounds__(256) pack_features_kernel(int P, const float *__restrict__ colors, float4 *__restrict__ feat) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    feat[i] = make_float4(colors[3 * (size_t)i], colors[3 * (size_t)i + 1], colors[3 * (size_t)i + 2], 0.f);
}

}  // namespace

#ifdef D3GA_BLEND_TRACE
extern "C" int d3ga_debug_bwd_trace(unsigned long long *out, int ctas) {
    return cudaMemcpyFromSymbol(out, g_bwd_trace, sizeof(unsigned long long) * 32 * (size_t)ctas) == cudaSuccess ? 0 : -2;
}
extern "C" int d3ga_debug_fwd_trace(unsigned long long *out, int ctas) {
    return cudaMemcpyFromSymbol(out, g_fwd_trace, sizeof(unsigned long long) * 32 * (size_t)ctas) == cudaSuccess ? 0 : -2;
}
#endif
#ifdef D3GA_BLEND_STATS
extern "C" int d3ga_debug_blend_stats(unsigned long long *out, int reset) {
    if (cudaMemcpyFromSymbol(out, g_blend_stats, sizeof(g_blend_stats)) != cudaSuccess) return -2;
    if (reset) {
        unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (cudaMemcpyToSymbol(g_blend_stats, z, sizeof(z)) != cudaSuccess) return -2;
    }
    return 0;
}
#endif

int launch_blend_forward(int P, int W, int H, const float *background, const float *colors, const uint8_t *geom,
                         const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, uint8_t *image,
                         const ImgLayout &IL, float *out_color, float *out_invdepth, cudaStream_t stream,
                         const float *background_b, const float *colors_b, float *out_color_b) {
    const bool pair = background_b || colors_b || out_color_b;
    if (pair && (!background_b || !colors_b || !out_color_b || out_invdepth)) return D3GA_EINVAL;
#ifdef FWD_BULK_STAGE
    if (pair) return D3GA_EINVAL;
#endif
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, num_tiles = gx * gy;
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    RasterHeader *hdr_rw = const_cast<RasterHeader *>(hdr);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const SplatRecord *splat = reinterpret_cast<const SplatRecord *>(geom + GL.splat);
    uint32_t *ballots = reinterpret_cast<uint32_t *>(image + IL.ballots);
    float *final_T = reinterpret_cast<float *>(image + IL.final_T);
    uint32_t *n_contrib = reinterpret_cast<uint32_t *>(image + IL.n_contrib);
    const uint32_t *order_fwd = reinterpret_cast<const uint32_t *>(image + IL.order_fwd);
    uint32_t *order_bwd = reinterpret_cast<uint32_t *>(image + IL.order_bwd);
    uint32_t *tile_nproc = reinterpret_cast<uint32_t *>(image + IL.tile_nproc);
    const size_t smem = (pair ? sizeof(FwdWarpSmemT<4>) : sizeof(FwdWarpSmemT<3>)) * FWD_WARPS;
    static std::atomic<uint64_t> attr_devices{0};
    const int arc = once_per_device(attr_devices, [&]() -> int {
        const int s3 = (int)(sizeof(FwdWarpSmemT<3>) * FWD_WARPS);
        D3GA_CUDA(cudaFuncSetAttribute(blend_forward_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s3));
        D3GA_CUDA(cudaFuncSetAttribute(blend_forward_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s3));
#ifndef FWD_BULK_STAGE
        D3GA_CUDA(cudaFuncSetAttribute(blend_forward_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(sizeof(FwdWarpSmemT<4>) * FWD_WARPS)));
#endif
        return D3GA_OK;
    });
    if (arc) return arc;
    // stage 1 packed the colours it was given into the geometry workspace; a render with OTHER colours on the same
    // geometry (the silhouette pass of a training frame) packs its own into its image workspace
    const float4 *feat = reinterpret_cast<const float4 *>(geom + GL.feat);
    if (colors != recall_packed_colors(geom)) {
        float4 *own = reinterpret_cast<float4 *>(image + IL.feat);
        feat = own;
        if (P > 0) {
            ProfScope prof(K_PACK_FEAT, stream);
            pack_features_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, colors, own);
            D3GA_LAUNCH_CHECK();
        }
    }
    float4 *feat_b = reinterpret_cast<float4 *>(image + IL.feat_b);
    if (pair && P > 0) {
        ProfScope prof(K_PACK_FEAT, stream);
        pack_features_kernel<<<(P + 255) / 256, 256, 0, stream>>>(P, colors_b, feat_b);
        D3GA_LAUNCH_CHECK();
    }
    ProfScope prof(K_BLEND_FWD, stream);
    {
        D3GA_CUDA(cudaMemsetAsync(tile_nproc, 0, sizeof(uint32_t) * (size_t)num_tiles, stream));
        if (out_invdepth)
            blend_forward_kernel<true, false><<<num_tiles, FWD_WARPS * 32, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, splat, feat, xy, co, background, final_T, n_contrib, out_color, out_invdepth,
                hdr_rw, gx, num_tiles, order_fwd, tile_nproc, order_bwd, ballots, nullptr, nullptr, nullptr);
#ifndef FWD_BULK_STAGE
        else if (pair)
            blend_forward_kernel<false, true><<<num_tiles, FWD_WARPS * 32, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, splat, feat, xy, co, background, final_T, n_contrib, out_color, nullptr,
                hdr_rw, gx, num_tiles, order_fwd, tile_nproc, order_bwd, ballots, feat_b, background_b, out_color_b);
#endif
        else
            blend_forward_kernel<false, false><<<num_tiles, FWD_WARPS * 32, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, splat, feat, xy, co, background, final_T, n_contrib, out_color, nullptr,
                hdr_rw, gx, num_tiles, order_fwd, tile_nproc, order_bwd, ballots, nullptr, nullptr, nullptr);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

int launch_blend_backward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                          const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, const uint8_t *image,
                          const ImgLayout &IL, const float *dL_dpix, const float *dL_dinvdepth_pix, float *grad_acc,
                          cudaStream_t stream, const float *background_b, const float *dL_dpix_b) {
    const bool pair = background_b || dL_dpix_b;
    if (pair && (!background_b || !dL_dpix_b || dL_dinvdepth_pix)) return D3GA_EINVAL;
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE, num_tiles = gx * gy;
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom + GL.header);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *v0 = reinterpret_cast<const uint32_t *>(binning + BL.vals0);
    const uint32_t *v1 = reinterpret_cast<const uint32_t *>(binning + BL.vals1);
    const float2 *xy = reinterpret_cast<const float2 *>(geom + GL.xy);
    const float4 *co = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    const SplatRecord *splat = reinterpret_cast<const SplatRecord *>(geom + GL.splat);
    const uint32_t *ballots = reinterpret_cast<const uint32_t *>(image + IL.ballots);
    const float4 *feat = colors == recall_packed_colors(geom) ? reinterpret_cast<const float4 *>(geom + GL.feat)
                                                              : reinterpret_cast<const float4 *>(image + IL.feat);
    const float *final_T = reinterpret_cast<const float *>(image + IL.final_T);
    const uint32_t *n_contrib = reinterpret_cast<const uint32_t *>(image + IL.n_contrib);
    const uint32_t *order_bwd = reinterpret_cast<const uint32_t *>(image + IL.order_bwd);
    const size_t smem = (pair ? sizeof(BwdWarpSmemT<4>) : sizeof(BwdWarpSmemT<3>)) * (TILE_PIX / 32);
    static std::atomic<uint64_t> attr_devices{0};
    const int arc = once_per_device(attr_devices, [&]() -> int {
        const int s3 = (int)(sizeof(BwdWarpSmemT<3>) * (TILE_PIX / 32));
        D3GA_CUDA(cudaFuncSetAttribute(blend_backward_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s3));
        D3GA_CUDA(cudaFuncSetAttribute(blend_backward_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, s3));
        D3GA_CUDA(cudaFuncSetAttribute(blend_backward_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(sizeof(BwdWarpSmemT<4>) * (TILE_PIX / 32))));
        return D3GA_OK;
    });
    if (arc) return arc;
    ProfScope prof(K_BLEND_BWD, stream);
    {
        const float4 *feat_b = reinterpret_cast<const float4 *>(image + IL.feat_b);   // packed by the pair forward
        if (dL_dinvdepth_pix)
            blend_backward_kernel<true, false><<<num_tiles, TILE_PIX, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, background, splat, feat, xy, co, final_T, n_contrib, dL_dpix,
                dL_dinvdepth_pix, grad_acc, gx, order_bwd, ballots, nullptr, nullptr, nullptr);
        else if (pair)
            blend_backward_kernel<false, true><<<num_tiles, TILE_PIX, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, background, splat, feat, xy, co, final_T, n_contrib, dL_dpix, nullptr,
                grad_acc, gx, order_bwd, ballots, feat_b, background_b, dL_dpix_b);
        else
            blend_backward_kernel<false, false><<<num_tiles, TILE_PIX, smem, stream>>>(
                hdr, ranges, v0, v1, W, H, background, splat, feat, xy, co, final_T, n_contrib, dL_dpix, nullptr,
                grad_acc, gx, order_bwd, ballots, nullptr, nullptr, nullptr);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
