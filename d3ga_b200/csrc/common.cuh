// common.cuh -- shared declarations of libd3ga_b200 (sm_100a).  Host + device.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <stddef.h>
#include <stdint.h>

#include "../../include/d3ga_b200.h"

namespace d3ga {

constexpr int TILE = 16;           // 16x16 pixel tiles (upstream BLOCK_X / BLOCK_Y)
constexpr int TILE_PIX = TILE * TILE;
constexpr int GEO_THREADS = 256;   // Gaussians per CTA in the per-Gaussian kernels
constexpr int SORT_THREADS = 512;
constexpr int SORT_ITEMS = 8;      // keys per thread per onesweep tile
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;  // 4096 pairs = 48 KB of payload per CTA
constexpr int SORT_MAX_PASSES = 8;
constexpr uint32_t MAX_DUPLICATES = (1u << 31) - 1;
// (gx+1)(gy+1) up to which the per-tile counts come from the rect difference grid (48 KB of shared memory)
constexpr int TILE_GRID_CELLS_MAX = 12288;
inline bool tile_plan_from_grid(int gx, int gy) { return (gx + 1) * (gy + 1) <= TILE_GRID_CELLS_MAX; }  // an inclusive look-back word keeps 1 flag bit

// ---- workspace layouts ------------------------------------------------------------------------------------------
struct RasterHeader {          // first 256 bytes of the geom workspace; zeroed at the start of every forward
    uint32_t num_rendered;     // D: total (tile, Gaussian) duplicates; may exceed capacity
    uint32_t overflow;         // 1 when D > capacity (later stages then do nothing)
    uint32_t geo_ticket;       // dynamic CTA id of the fused geometry kernel (look-back needs in-order ids)
    uint32_t final_sel;        // which key/value buffer (0/1) holds the sorted list
    uint32_t hist_done;        // CTAs of the histogram kernel that have finished
    uint32_t sort_bits;        // 32 + bits(tile id)
    uint32_t num_passes;
    uint32_t ranges_done;      // CTAs of the ranges kernel that have finished (last one orders the tiles)
    uint32_t pass_skip[SORT_MAX_PASSES];   // 1 = every key has the same digit in this pass
    uint32_t pass_src[SORT_MAX_PASSES];    // buffer (0/1) a pass reads from
    uint32_t pass_ticket[SORT_MAX_PASSES]; // dynamic tile id per pass
    uint32_t blend_done;       // CTAs of the blend forward that have finished (last one orders the backward)
    // plan of the 32-bit depth sort over the Gaussians (depth-first binning)
    uint32_t dep_final_sel;    // which depth key/value buffer (0/1) holds the depth-sorted Gaussian list
    uint32_t dep_hist_done;
    uint32_t dep_culled;       // Gaussians with no tile (their key is 0xffffffff)
    uint32_t dep_skip[4];
    uint32_t dep_src[4];
    uint32_t dep_ticket[4];
    // direct tile scatter (raster_direct.cu)
    uint64_t direct_total;     // duplicates counted so far by the count rows
    uint32_t direct_done;      // rows of direct_count_kernel that have finished (the last one publishes num_rendered)
    uint32_t pad1[64 - 9 - 3 * SORT_MAX_PASSES - 3 - 12 - 3];
};
static_assert(offsetof(RasterHeader, direct_total) % 8 == 0, "64-bit counter must be aligned");
static_assert(sizeof(RasterHeader) == 256, "header is 256 bytes");

// The digit counts of the depth keys are kept in DEP_HIST_COPIES copies (a CTA adds to copy blockIdx % copies): the top
// digits of the keys take two or three values, and thousands of CTAs adding to the same few words of ONE copy serialise in
// the L2 (measured at 1 M Gaussians: preprocess 36 us with 8 or more copies, 58 us with one).
#ifndef D3GA_DEP_HIST_COPIES
#define D3GA_DEP_HIST_COPIES 8
#endif
constexpr int DEP_HIST_COPIES = D3GA_DEP_HIST_COPIES;
struct GeomLayout {
    uint64_t header, depth, xy, conic_opacity, rgb, cov3D, tiles, offsets, clamped, scan_state;
    uint64_t rect_pack, dkeys, dvals, dep_hist, dep_bins, dep_state, tile_diff, total;   // depth-first binning
    uint64_t ordered;    // direct tile scatter: the Gaussians' packed records in depth order (16 B each)
    uint64_t grec;       // ... and in Gaussian order, as the preprocess wrote them
    uint64_t splat;      // what the blend kernels gather per list entry: 2 x float4 per Gaussian (SplatRecord below)
    uint64_t feat;       // ... and the colours stage 1 was given, packed (r, g, b, 0) -- see ImgLayout::feat
    uint64_t p_pad;
};
struct BinLayout {
    uint64_t keys0, keys1, vals0, vals1, hist, bins, tile_state, total;
    uint64_t tile_state_bytes;
};
struct ImgLayout {
    uint64_t ranges, final_T, n_contrib, order_fwd, order_bwd, tile_nproc, ballots, feat, feat_b, total;
};
// The blend kernels' per-Gaussian gather: ONE 32-byte sector, two 16-byte loads (before: xy 8 B + conic/opacity 16 B +
// depth 4 B from three arrays).  The conic is pre-scaled to the log2 domain (log2 G = A dx^2 + B dx dy + C dy^2) and the
// footprint bound of the per-block cull comes with it, so no logarithm is evaluated per (warp, list entry).
//   g0 = (x, y, lim2, opacity)      g1 = (A, B, C, 1 / depth)
//   lim2: a pixel block can hold a pixel with alpha >= 1/255 only if max over the block of (A dx^2 + B dx dy + C dy^2)
//         >= lim2;  +inf = never (opacity below 1/255), -inf = never culled (degenerate conic)
struct SplatRecord {
    float4 g0, g1;
};
static_assert(sizeof(SplatRecord) == 32, "one sector per Gaussian");
// Survivor ballots of the forward blend, replayed by the backward: one word per (32-entry chunk of a tile's list, warp).
// Tile t's chunks start at word row (range.x / 32 + t): consecutive tiles never overlap (ceil(n / 32) <= n / 32 + 1).
__host__ __device__ inline uint64_t ballot_rows(uint64_t capacity, uint64_t tiles) { return capacity / 32 + tiles + 1; }

__host__ __device__ inline uint64_t align256(uint64_t x) { return (x + 255) & ~uint64_t(255); }

inline GeomLayout geom_layout(int P) {
    GeomLayout L;
    uint64_t o = 0, p = (uint64_t)P;
    L.header = o;        o += 256;
    L.depth = o;         o = align256(o + 4 * p);
    L.xy = o;            o = align256(o + 8 * p);
    L.conic_opacity = o; o = align256(o + 16 * p);
    L.rgb = o;           o = align256(o + 12 * p);
    L.cov3D = o;         o = align256(o + 24 * p);
    L.tiles = o;         o = align256(o + 4 * p);
    L.offsets = o;       o = align256(o + 4 * p);
    L.clamped = o;       o = align256(o + 3 * p);
    L.scan_state = o;    o = align256(o + 8 * ((p + GEO_THREADS - 1) / GEO_THREADS + 1));
    L.p_pad = align256(p < 1 ? 1 : p);   // element stride between the two halves of the depth-sort ping-pong
    L.rect_pack = o;     o = align256(o + 8 * p);
    L.dkeys = o;         o = align256(o + 8 * L.p_pad);
    L.dvals = o;         o = align256(o + 8 * L.p_pad);
    L.dep_hist = o;      o = align256(o + 4 * 256 * 4 * DEP_HIST_COPIES);
    L.dep_bins = o;      o = align256(o + 4 * 256 * 4);
    L.dep_state = o;     o = align256(o + 4ull * 256 * 4 * ((p + SORT_TILE - 1) / SORT_TILE + 1));
    L.tile_diff = o;     o = align256(o + 4ull * TILE_GRID_CELLS_MAX);   // 2-D difference grid of the tile rects
    L.ordered = o;       o = align256(o + 16 * L.p_pad);
    L.grec = o;          o = align256(o + 16 * L.p_pad);
    L.splat = o;         o = align256(o + 32 * p);
    L.feat = o;          o = align256(o + 16 * p);
    L.total = o;
    return L;
}
inline uint64_t sort_tiles(uint64_t capacity) { return (capacity + SORT_TILE - 1) / SORT_TILE; }
inline BinLayout bin_layout(uint64_t capacity) {
    BinLayout L;
    uint64_t o = 0, c = capacity < 1 ? 1 : capacity;
    L.keys0 = o; o = align256(o + 8 * c);
    L.keys1 = o; o = align256(o + 8 * c);
    L.vals0 = o; o = align256(o + 4 * c);
    L.vals1 = o; o = align256(o + 4 * c);
    L.hist = o;  o = align256(o + 4 * 256 * SORT_MAX_PASSES);   // digit counts per pass
    L.bins = o;  o = align256(o + 4 * 256 * SORT_MAX_PASSES);   // exclusive prefix per pass
    L.tile_state = o;
    L.tile_state_bytes = 4ull * 256 * sort_tiles(c) * SORT_MAX_PASSES;
    o = align256(o + L.tile_state_bytes);
    L.total = o;
    return L;
}
inline ImgLayout img_layout(int P, int W, int H, uint64_t capacity) {
    ImgLayout L;
    uint64_t o = 0, tiles = (uint64_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE), px = (uint64_t)W * H;
    L.ranges = o;    o = align256(o + 8 * tiles);
    L.final_T = o;   o = align256(o + 4 * px);
    L.n_contrib = o; o = align256(o + 4 * px);
    L.order_fwd = o;  o = align256(o + 4 * tiles);   // tiles, longest list first (blend forward launch order)
    L.order_bwd = o;  o = align256(o + 4 * tiles);   // tiles, deepest contributor first (blend backward)
    L.tile_nproc = o; o = align256(o + 4 * tiles);   // max over the tile's pixels of n_contrib
    L.ballots = o;    o = align256(o + 4ull * 8 * ballot_rows(capacity, tiles));
    // this render's colours as one aligned 16-byte vector per Gaussian (r, g, b, 0), when they are not the colours stage 1
    // packed into the geometry workspace: the second render of a training frame shares that workspace but blends
    // other colours (the blend launchers pick by the colour pointer: recall_packed_colors)
    L.feat = o;       o = align256(o + 16ull * (uint64_t)(P < 1 ? 1 : P));
    // the second image's colours of a PAIR render (d3ga_raster_forward_render_pair), packed the same way
    L.feat_b = o;     o = align256(o + 16ull * (uint64_t)(P < 1 ? 1 : P));
    L.total = o;
    return L;
}

// Packed per-Gaussian gradient accumulator of the blend backward: three 16-byte vectors per Gaussian so one
// (tile, Gaussian) pair costs three vector reductions (REDG.E.ADD.F32x4) instead of ten scalar atomics.
//   v0 = (dmean2D.x, dmean2D.y, dconic.xx, dconic.xy)   v1 = (dconic.yy, dopacity, dcolor.r, dcolor.g)
//   v2 = (dcolor.b, dinvdepth, 0, 0)
constexpr int GRAD_ACC_FLOATS = 12;

// ---- host-side helpers -----------------------------------------------------------------------------------------
enum KernelId {
    K_GEOMETRY = 0, K_SORT_HIST, K_SORT_PASS, K_RANGES, K_BLEND_FWD, K_BLEND_BWD, K_PRE_BWD, K_DEFORM_FWD, K_DEFORM_BWD,
    K_BARY, K_PREPROCESS, K_EMIT, K_DEPTH_HIST, K_DEPTH_PASS, K_FEM_FWD, K_FEM_BWD, K_SSIM_FWD, K_SSIM_BWD,
    K_LBS_SKELETON, K_LBS_SKIN_FWD, K_LBS_SKIN_BWD, K_LBS_UNSKIN, K_LBS_DENSE_FWD, K_LBS_DENSE_BWD, K_RECT_GRID, K_TILE_PLAN,
    K_DIRECT_COUNT, K_DIRECT_SEGSUM, K_DIRECT_PLAN, K_DIRECT_BASE, K_DIRECT_SCATTER, K_PACK_FEAT, K_DEFORM_SELECT, K_COUNT
};
// Optional per-kernel CUDA-event timing (d3ga_profile_* in the C-ABI; bench.py's roofline figures come from it).
// A no-op unless profiling is enabled.
struct ProfScope {
    int id;
    cudaStream_t stream;
    void *slot;
    ProfScope(int id, cudaStream_t stream);
    ~ProfScope();
};
void count_launch(int n = 1);
int set_cuda_error(cudaError_t e);
#define D3GA_CUDA(expr)                                         \
    do {                                                        \
        cudaError_t _e = (expr);                                \
        if (_e != cudaSuccess) return d3ga::set_cuda_error(_e); \
    } while (0)
#define D3GA_LAUNCH_CHECK()                                     \
    do {                                                        \
        d3ga::count_launch();                                   \
        cudaError_t _e = cudaPeekAtLastError();                 \
        if (_e != cudaSuccess) return d3ga::set_cuda_error(_e); \
    } while (0)

// Function attributes (the opt-in to more than 48 KB of dynamic shared memory) belong to the device that is current when
// they are set, so a launcher applies them the first time it runs on EACH device: `mask` has one bit per device
// ordinal (ordinals >= 64 simply set them on every call).  f() returns a D3GA_* code.
template <typename F>
inline int once_per_device(std::atomic<uint64_t> &mask, F &&f) {
    int dev = 0;
    D3GA_CUDA(cudaGetDevice(&dev));
    const uint64_t bit = dev >= 0 && dev < 64 ? (1ull << dev) : 0ull;
    if (bit && (mask.load(std::memory_order_acquire) & bit)) return D3GA_OK;
    const int rc = f();
    if (rc == D3GA_OK && bit) mask.fetch_or(bit, std::memory_order_release);
    return rc;
}
int device_sm_count();   // multiprocessors of the current device (cached per device)

// stage launchers (defined in the per-stage .cu files)
int launch_geometry(int P, int sh_degree, int sh_coeffs, const float *means3D, const float *scales,
                    float scale_modifier, const float *rotations, const float *opacities, const float *shs,
                    const float *cov3D_precomp, const float *colors_precomp, const float *viewmatrix,
                    const float *projmatrix, const float *campos, int W, int H, float tanfovx, float tanfovy,
                    int antialiasing, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                    uint64_t capacity, int32_t *radii, cudaStream_t stream);
int launch_mark_visible(int P, const float *means3D, const float *viewmatrix, uint8_t *present, cudaStream_t stream);
int launch_sort_and_ranges(int P, uint8_t *geom, const GeomLayout &GL, int gx, int gy, uint8_t *binning, const BinLayout &BL, uint64_t capacity, int sort_bits,
                           uint8_t *image, const ImgLayout &IL, int num_tiles, cudaStream_t stream);
int launch_depth_sort(int P, uint8_t *geom, const GeomLayout &GL, cudaStream_t stream);
enum SortMode { SORT_DEPTH_FIRST = 0, SORT_GLOBAL = 1, SORT_DIRECT = 3 };
int sort_mode();   // D3GA_SORT_MODE = direct (default) | depthfirst (its fallback) | global (upstream's order: cross-check)
// the two modes that sort the Gaussians by depth first; `direct` falls back to `depthfirst` where its plan does not apply
inline bool depth_first_mode(int mode) { return mode == SORT_DEPTH_FIRST || mode == SORT_DIRECT; }
// Direct tile scatter (raster_direct.cu): rows of `chunk` depth-ordered Gaussians, one private tile histogram per row.
constexpr int DIRECT_MAX_TILES = 49152;   // the plan CTA keeps 4 B per tile in shared memory
constexpr int DIRECT_MAX_CELLS = 53248;   // (gx+1)(gy+1): the count CTA's difference grid, 4 B per cell (208 KB)
constexpr int DIRECT_MAX_CHUNK = 8192;    // a (row, tile) count must fit 16 bits
struct DirectPlan {
    int ok;                    // 0: not applicable (mode, tile grid too large, capacity too small for the scratch)
    uint32_t chunk, rows, tl_pad, nseg, rows_per_seg, colblocks;
    uint32_t bands, band_rows, band_tiles_pad;   // the scatter's split of the image into bands of tile rows
    uint64_t seg_off, tot_off, start_off, ticket_off, before_off;   // byte offsets inside the unused key buffer
};
DirectPlan direct_plan(int P, int gx, int gy, uint64_t capacity);   // stage 1 only: reads the views-in-flight hint
// the plan stage 1 used on this geometry workspace (false: no stage 1 has run on it with these sizes)
void remember_direct_plan(const void *geom, int P, int gx, int gy, uint64_t capacity, const DirectPlan &pl,
                          const void *packed_colors);
bool recall_direct_plan(const void *geom, int P, int gx, int gy, uint64_t capacity, DirectPlan *pl);
// the colour array stage 1 packed into the geometry workspace's `feat` section (null: none / unknown workspace)
const void *recall_packed_colors(const void *geom);
void set_views_in_flight(int n);
int launch_direct_count(int P, int gx, int gy, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                        uint64_t capacity, const DirectPlan &pl, cudaStream_t stream);
int launch_sorted_keys(int num_tiles, const uint8_t *geom, const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL,
                       const uint8_t *image, const ImgLayout &IL, uint64_t *keys_out, cudaStream_t stream);
int launch_copy_sorted_keys(const uint8_t *geom, const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL,
                            uint64_t capacity, uint64_t *keys_out, cudaStream_t stream);
int launch_direct_scatter(int P, int gx, int gy, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                          uint64_t capacity, int sort_bits, uint8_t *image, const ImgLayout &IL, const DirectPlan &pl,
                          cudaStream_t stream);
int launch_blend_forward(int P, int W, int H, const float *background, const float *colors, const uint8_t *geom,
                         const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, uint8_t *image,
                         const ImgLayout &IL, float *out_color, float *out_invdepth, cudaStream_t stream,
                         const float *background_b = nullptr, const float *colors_b = nullptr,
                         float *out_color_b = nullptr);   // the three together: a PAIR render (no inverse depth)
int launch_blend_backward(int W, int H, const float *background, const float *colors, const uint8_t *geom,
                          const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL, const uint8_t *image,
                          const ImgLayout &IL, const float *dL_dpix, const float *dL_dinvdepth_pix, float *grad_acc,
                          cudaStream_t stream, const float *background_b = nullptr, const float *dL_dpix_b = nullptr);
int launch_preprocess_backward(int P, int sh_degree, int sh_coeffs, int W, int H, const float *means3D,
                               const float *opacities, const float *scales, float scale_modifier,
                               const float *rotations, const float *cov3D_precomp, const float *shs,
                               const float *viewmatrix, const float *projmatrix, const float *campos, float tanfovx,
                               float tanfovy, int antialiasing, const int32_t *radii, const uint8_t *geom,
                               const GeomLayout &GL, const float *grad_acc, int has_invdepth_grad,
                               float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors, float *dL_dmeans3D,
                               float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                               cudaStream_t stream);

#ifdef __CUDACC__
// ---- device helpers ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
// Lanes holding the same 8-bit digit.  Eight ballots + logic instead of MATCH.ANY: on sm_100 the match instruction is
// issued through the MIO queue and throttles a ranking loop (ncu: stall_mio on every MATCH.ANY), ballots are plain
// ALU-side votes.
__device__ __forceinline__ uint32_t match_digit8(uint32_t d) {
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool bit = (d >> b) & 1u;
        const uint32_t m = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? m : ~m;
    }
    return peers;
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ uint64_t ld_volatile_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_volatile_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Longest-processing-time-first launch order: one CTA of 256 threads counting-sorts the tiles by key, descending
// (256 buckets of width 32, saturating; order inside a bucket is arbitrary).  Two passes over the keys with the loads
// batched eight deep (a single CTA is latency-bound on them).  s_bins needs 256 words.
template <typename KeyFn>
__device__ __forceinline__ void order_tiles_descending(int num_tiles, KeyFn key, uint32_t *order, uint32_t *s_bins) {
    const int tid = threadIdx.x;
    __shared__ uint32_t s_w[8];
    s_bins[tid] = 0;
    __syncthreads();
    for (int t0 = 0; t0 < num_tiles; t0 += 256 * 8) {
        uint32_t k[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int t = t0 + u * 256 + tid; k[u] = t < num_tiles ? key(t) : 0u; }
#pragma unroll
        for (int u = 0; u < 8; u++)
            if (t0 + u * 256 + tid < num_tiles) atomicAdd(&s_bins[255u - min(255u, k[u] >> 5)], 1u);
    }
    __syncthreads();
    const uint32_t lane = tid & 31, warp = tid >> 5;
    const uint32_t c = s_bins[tid];
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w = 0; w < warp; w++) off += s_w[w];
    s_bins[tid] = off + incl - c;
    __syncthreads();
    for (int t0 = 0; t0 < num_tiles; t0 += 256 * 8) {
        uint32_t k[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int t = t0 + u * 256 + tid; k[u] = t < num_tiles ? key(t) : 0u; }
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const int t = t0 + u * 256 + tid;
            if (t < num_tiles) order[atomicAdd(&s_bins[255u - min(255u, k[u] >> 5)], 1u)] = (uint32_t)t;
        }
    }
}

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_v2(float *addr, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1,%2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}
#endif

// exclusive scan of one value per thread over the FIRST 256 threads of a CTA (every thread of the CTA must call it:
// it contains barriers); s_tmp needs 8 words
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t *s_tmp, uint32_t *total) {
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    __syncthreads();   // s_tmp may still be in use by a previous call
    if (lane == 31 && warp < 8) s_tmp[warp] = incl;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t x = s_tmp[w];
        if (w < (int)warp) off += x;
        tot += x;
    }
    if (total) *total = tot;
    return off + incl - v;
}


// The plan of the 32-bit depth sort of the Gaussians, from the digit counts `ghist` [DEP_HIST_COPIES][4][256] of their
// keys: bin offsets per pass, which passes run and which buffer each reads.  Culled Gaussians carry the key 0xffffffff:
// they are counted (they move in every executed pass) but do not stop a pass from being skipped -- a pass in which every
// VISIBLE key has the same digit is the identity for the visible keys.  One CTA of 1024 threads, thread (pass, digit); a
// thread's loads of the copies are all in flight together (the CTA is alone on the GPU: latency is all it costs).
__device__ __forceinline__ void depth_sort_plan(RasterHeader *hdr, const uint32_t *ghist, uint32_t *bins, uint32_t n) {
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_one[4];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, p = tid >> 8, d = tid & 255u;
    if (tid < 4) s_one[tid] = 0;
    uint32_t part[DEP_HIST_COPIES];
#pragma unroll
    for (int r = 0; r < DEP_HIST_COPIES; r++) part[r] = __ldcg(ghist + r * 1024 + tid);
    const uint32_t n_culled = *((volatile uint32_t *)&hdr->dep_culled);
    uint32_t c = 0;
#pragma unroll
    for (int r = 0; r < DEP_HIST_COPIES; r++) c += part[r];
    const uint32_t n_visible = n - n_culled;
    uint32_t incl = c;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, s);
        if (lane >= (uint32_t)s) incl += v;
    }
    __syncthreads();
    if (lane == 31) s_w[warp] = incl;
    if (n_visible == 0 || c - (d == 255u ? n_culled : 0u) == n_visible) s_one[p] = 1u;   // every visible key in this bin
    __syncthreads();
    uint32_t off = 0;
    for (uint32_t w = warp & ~7u; w < warp; w++) off += s_w[w];
    bins[tid] = off + incl - c;
    if (tid == 0) {
        uint32_t src = 0;
        for (int q = 0; q < 4; q++) {
            hdr->dep_skip[q] = s_one[q];
            hdr->dep_src[q] = src;
            if (!s_one[q]) src ^= 1u;
        }
        hdr->dep_final_sel = src;
    }
}

// ---- [N,K] float rows of 32 consecutive Gaussians, moved as coalesced requests ------------------------------------
// The warp's 32 consecutive rows of K floats lie contiguous in memory: they are read as K coalesced requests of four
// sectors each and every lane takes its row out of shared memory (stride K).  A scalar load of ONE column of a [N,K]
// array touches up to 9 K sectors per request; ncu showed the deformation forward's LSU queue full (lg_throttle 15
// stalls per issue, 27 M sector accesses for 120 MB).  All 32 lanes call (rows past the end read as 0 / are not stored).
constexpr int STAGE_FLOATS = 32 * 9;
// Two phases, so that a kernel can issue the global loads of ALL its streams before it waits for the first one.
template <int K>
__device__ __forceinline__ void warp_rows_fetch(const float *__restrict__ base, size_t row0, size_t rows, uint32_t lane,
                                                float (&raw)[K]) {
    const size_t first = row0 * K, limit = rows * K;
#pragma unroll
    for (int j = 0; j < K; j++) {
        const size_t i = first + (size_t)(j * 32) + lane;
        raw[j] = i < limit ? __ldg(base + i) : 0.f;
    }
}
template <int K>
__device__ __forceinline__ void warp_rows_take(float *s_w, uint32_t lane, const float (&raw)[K], float (&out)[K]) {
#pragma unroll
    for (int j = 0; j < K; j++) s_w[j * 32 + (int)lane] = raw[j];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < K; j++) out[j] = s_w[(int)lane * K + j];
    __syncwarp();
}
template <int K>
__device__ __forceinline__ void warp_rows_store(float *__restrict__ base, size_t row0, size_t rows, float *s_w,
                                                uint32_t lane, const float (&in)[K]) {
    const size_t first = row0 * K, limit = rows * K;
#pragma unroll
    for (int j = 0; j < K; j++) s_w[(int)lane * K + j] = in[j];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < K; j++) {
        const size_t i = first + (size_t)(j * 32) + lane;
        if (i < limit) base[i] = s_w[j * 32 + (int)lane];
    }
    __syncwarp();
}

}  // namespace d3ga
