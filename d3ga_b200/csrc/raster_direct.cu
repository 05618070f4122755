// raster_direct.cu -- rows R2-R5 of SURVEY.md section 8 (scan, key duplication, sort, per-tile ranges) WITHOUT sorting
// the duplicates: a stable counting scatter by tile id, fed by the depth-ordered Gaussian list of the depth-first binning.
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::SortPairs + identifyTileRanges inside
// diff_gaussian_rasterization._C.rasterize_gaussians (reached from renderer.py:130-141); the sorted (key, Gaussian) list,
// the ranges and the duplicate count it leaves in the workspaces are bit-identical to theirs.
//
// Once the N Gaussians are in depth order (raster_sort.cu: launch_depth_sort), the sorted duplicate list is fully
// determined by COUNTS: the slot of duplicate (Gaussian g, tile t) is
//     start[t]  +  #duplicates of t among the Gaussians before g in depth order.
// The depth-ordered list is cut into rows of `chunk` Gaussians, one CTA per row:
//   1. direct_count_kernel   : per-tile counts of the row (a 2-D difference grid in shared memory: four corner updates per
//                              rect, then prefix sums) -> M[row][tile] (u16); also leaves the row's Gaussians as packed
//                              records (rect, extent, id, depth bits) in depth order so the scatter reads them coalesced,
//                              and adds the row's total to D (num_rendered, overflow).
//   2. direct_segsum_kernel  : column sums of M over segments of rows; the last CTA of a column block turns them into
//                              exclusive segment offsets and the per-tile totals.
//   3. direct_plan_kernel    : one CTA: exclusive scan of the totals in tile order = ranges (a tile without duplicates
//                              keeps (0, 0), like upstream's zero-initialised ranges) and the blend's tile launch order.
//   4. direct_base_kernel    : segment offsets += start[tile]; before[row][tile] = exclusive prefix of M inside the segment.
//   5. direct_scatter_kernel : the row keeps pos[tile] (next free slot) and mask[tile] in shared memory.  Per batch of 32
//                              rect pieces three walks: (A) every owner ORs its bit into mask[t] of its tiles; (B) the
//                              slot of (owner, t) is pos[t] + popc(mask[t] below the owner) -- owners that share a tile
//                              are ordered by position in the batch = by depth, whatever order the walks reach the tile
//                              in -- and the Gaussian id is stored straight into its final slot; (C) the highest owner of
//                              mask[t] advances pos[t] by popc(mask[t]) and clears the mask.  (Details above the kernel.)
// Per duplicate that is three cheap visits and one 4-byte store, against an emission (12 B store) plus two radix passes
// (each 12 B load + MATCH-based ranking + 12 B store) before; a row is independent of every other row and nothing looks
// back at a predecessor.  The sorted keys are not stored (the blend does not read them; sorted_keys_kernel rebuilds them
// on request).  Algorithmic bytes: 36 N (gathers + records) in (1), 16 N + 4 D in (5); the count rows and their prefixes
// (6 B x rows x tiles) are scratch traffic.
// Scratch (M, segment sums, prefixes) lives in the second halves of the binning workspace's key / value ping-pong, which
// this path does not use.  The row length is chosen so that the rows fill the GPU once -- or 1/n of it when the caller
// keeps n views in flight (d3ga_set_views_in_flight) -- at 8 B x tiles of shared memory per row; larger images are cut into
// bands of tile rows.  The path is not taken (the radix sort runs instead) for tile grids beyond DIRECT_MAX_TILES or when
// the scratch does not fit.
#include <atomic>
#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace d3ga {

namespace {

constexpr int COL_THREADS = 256;       // tile pairs per CTA in the column kernels
constexpr int DIRECT_MAX_SEGS = 32;
constexpr uint64_t SCATTER_SMEM_BUDGET = 36 * 1024;

struct DirectParams {
    RasterHeader *hdr;
    int P, gx;
    uint32_t P_pad, tl_pad, chunk, rows, rows_per_seg;
    uint64_t capacity;
    const uint32_t *dvals;     // [2][P_pad] values of the depth sort (Gaussian ids in depth order)
    const uint4 *grec;         // [P] per-Gaussian records written by the preprocess (same fields as rec)
    uint4 *rec;                // [P_pad] depth-ordered records: x0 | y0 << 16, width | height << 16, Gaussian id, depth bits
    uint16_t *M;               // [rows][tl_pad] duplicates of tile t in row r
    const uint16_t *before;    // [rows][tl_pad] duplicates of tile t in the rows before r inside r's segment
    const uint32_t *seg;       // [nseg][tl_pad] slot of the segment's first duplicate of tile t
    uint32_t *vals;
};

// One CTA per row.  The per-tile counts of the row's Gaussians come from a 2-D difference grid in shared memory: +1 over a
// tile rect is four corner updates (one Gaussian per lane, every lane busy, whatever the rect sizes), and the prefix sums
// along x and then y of the grid are the counts.  The row's batches are dealt round-robin to the warps -- counting does not
// care about order.  Also leaves the row's records in depth order for the scatter and adds the row's total to D.
constexpr int COUNT_WARPS = 8;
__global__ void __launch_bounds__(COUNT_WARPS * 32) direct_count_kernel(const DirectParams p, int gy, int num_tiles) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ unsigned long long s_sum;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, row = blockIdx.x;
    const int gx = p.gx, pitch = gx + 1, cells = pitch * (gy + 1);
    int32_t *grid = reinterpret_cast<int32_t *>(smem);
    for (int i = (int)tid; i < cells; i += COUNT_WARPS * 32) grid[i] = 0;
    if (tid == 0) s_sum = 0;
    __syncthreads();
    const uint32_t *order = p.hdr->dep_final_sel ? p.dvals + p.P_pad : p.dvals;
    const uint32_t slot0 = row * p.chunk, nb = p.chunk / 32;
    // Gaussian ids run two batches ahead, their records one batch
    auto load_id = [&](uint32_t b) {
        const uint32_t s = slot0 + b * 32 + lane;
        return (b < nb && s < (uint32_t)p.P) ? order[s] : 0xffffffffu;
    };
    auto gather = [&](uint32_t g) { return g == 0xffffffffu ? make_uint4(0, 0, 0, 0) : p.grec[g]; };
    uint32_t id_nn = load_id(warp);
    uint4 rec_n = gather(id_nn);
    id_nn = load_id(warp + COUNT_WARPS);
    uint32_t sum = 0;
    for (uint32_t b = warp; b < nb; b += COUNT_WARPS) {
        const uint4 rec = rec_n;
        rec_n = gather(id_nn);
        id_nn = load_id(b + 2 * COUNT_WARPS);
        const uint32_t s = slot0 + b * 32 + lane;
        if (s < (uint32_t)p.P) p.rec[s] = rec;
        const uint32_t w = rec.y & 0xffffu, h = rec.y >> 16;
        if (w * h) {
            const int x0 = (int)(rec.x & 0xffffu), y0 = (int)(rec.x >> 16), x1 = x0 + (int)w, y1 = y0 + (int)h;
            atomicAdd(&grid[y0 * pitch + x0], 1);
            atomicAdd(&grid[y0 * pitch + x1], -1);
            atomicAdd(&grid[y1 * pitch + x0], -1);
            atomicAdd(&grid[y1 * pitch + x1], 1);
            sum += w * h;
        }
    }
    sum = __reduce_add_sync(0xffffffffu, sum);
    if (lane == 0 && sum) atomicAdd(&s_sum, (unsigned long long)sum);
    __syncthreads();
    for (int y = (int)tid; y < gy; y += COUNT_WARPS * 32) {   // prefix along x (odd pitches are bank-conflict free)
        int run = 0;
        for (int x = 0; x < gx; x++) { run += grid[y * pitch + x]; grid[y * pitch + x] = run; }
    }
    __syncthreads();
    uint16_t *Mrow = p.M + (size_t)row * p.tl_pad;
    for (int x = (int)tid; x < gx; x += COUNT_WARPS * 32) {   // prefix along y = the counts; neighbouring threads, neighbouring tiles
        int run = 0;
        for (int y = 0; y < gy; y++) { run += grid[y * pitch + x]; Mrow[y * gx + x] = (uint16_t)run; }
    }
    for (uint32_t t = (uint32_t)num_tiles + tid; t < p.tl_pad; t += COUNT_WARPS * 32) Mrow[t] = 0;
    if (tid == 0) {
        if (s_sum) atomicAdd(reinterpret_cast<unsigned long long *>(&p.hdr->direct_total), s_sum);
        __threadfence();
        if (atomicAdd(&p.hdr->direct_done, 1u) == p.rows - 1) {   // last row: the duplicate count is complete
            __threadfence();
            const unsigned long long total = atomicAdd(reinterpret_cast<unsigned long long *>(&p.hdr->direct_total), 0ull);
            p.hdr->num_rendered = total > 0xffffffffull ? 0xffffffffu : (uint32_t)total;
            p.hdr->overflow = total > p.capacity ? 1u : 0u;
        }
    }
}

// One CTA of SC_WARPS warps per row.  The image's tile rows are dealt to the warps in blocks of four (tile row y belongs to
// warp (y / 4) mod SC_WARPS), so the warps share pos[] / mask[] without ever touching the same word and need no
// synchronisation after the set-up, and a body that fills a third of the image still loads them evenly.
// Every warp runs over all the records of the row, keeps the pieces of the rects that fall into its own blocks (one ENTRY
// per rect and block: at most four tile rows) in a small FIFO in shared memory, and whenever 32 entries are queued
// processes them as one batch, one entry per lane:
//   - the entries are cut into STRIPS (one tile row each: first tile, width, owner lane) listed in shared memory, so that
//     the walks run over strips, one per lane: a strip is 1-4 tiles wide almost always, while whole rects span 1 to 70
//     tiles -- a lane per rect would leave two thirds of the warp idle;
//   - (A) every strip ORs its owner's bit into mask[t] of its tiles; (B) the slot of (owner, t) is pos[t] + popc(mask[t]
//     below the owner) -- owners that share a tile are ordered by queue position = by depth, whatever order the walks reach
//     the tile in -- and the value and the (tile | depth) key go straight into their final slots; (C) the highest owner of
//     mask[t] advances pos[t] by popc(mask[t]) and clears the mask.
#ifndef DIRECT_BLK_SHIFT
#define DIRECT_BLK_SHIFT 0
#endif
#ifndef DIRECT_SC_WARPS
#define DIRECT_SC_WARPS 4
#endif
constexpr int SC_WARPS = DIRECT_SC_WARPS, BLK_SHIFT = DIRECT_BLK_SHIFT, BLK_ROWS = 1 << BLK_SHIFT;
constexpr int QUEUE = 64, STRIP_CAP = 32 * BLK_ROWS, STRIP_ROUNDS = STRIP_CAP / 32;

struct ScatterWarp {
    uint4 queue[QUEUE];            // entries: x0 | y0 << 16, width | rows << 16, Gaussian id, depth bits
    uint32_t strips[STRIP_CAP];    // first tile | width << 16 | owner lane << 27
    uint32_t own[32];              // Gaussian id of the batch's lanes
};

#ifndef DIRECT_INLINE
#define DIRECT_INLINE __forceinline__
#endif
__device__ DIRECT_INLINE void scatter_batch(const DirectParams &p, ScatterWarp &sw, uint2 *pm, uint32_t head, uint32_t count,
                                              uint32_t lane) {
    const uint4 e = lane < count ? sw.queue[(head + lane) & (QUEUE - 1)] : make_uint4(0, 0, 0, 0);
    const uint32_t w = e.y & 0xffffu, rows = w ? (e.y >> 16) : 0u, gx = (uint32_t)p.gx;
    sw.own[lane] = e.z;
    uint32_t incl = rows;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    uint32_t se = ((e.x >> 16) * gx + (e.x & 0xffffu)) | (w << 16) | (lane << 27);
    for (uint32_t i = 0, o = incl - rows; i < rows; i++, o++, se += gx) sw.strips[o] = se;
    __syncwarp();
    // the batch's strips, one per lane and round, kept in registers for the three walks
    uint32_t t0[STRIP_ROUNDS], sn[STRIP_ROUNDS], src[STRIP_ROUNDS], wm[STRIP_ROUNDS];
#pragma unroll
    for (int r = 0; r < STRIP_ROUNDS; r++) {
        const uint32_t v = (uint32_t)r * 32 + lane < total ? sw.strips[r * 32 + lane] : 0u;
        t0[r] = v & 0xffffu; sn[r] = (v >> 16) & 0x7ffu; src[r] = v >> 27;
        wm[r] = (uint32_t)r * 32 < total ? __reduce_max_sync(0xffffffffu, sn[r]) : 0u;
    }
#pragma unroll
    for (int r = 0; r < STRIP_ROUNDS; r++)
        for (uint32_t k = 0; k < wm[r]; k++)
            if (k < sn[r]) atomicOr(&pm[t0[r] + k].y, 1u << src[r]);
    __syncwarp();
#pragma unroll
    for (int r = 0; r < STRIP_ROUNDS; r++) {
        if (wm[r] == 0) continue;
        const uint32_t id = sw.own[src[r]];
        const uint32_t lower = (1u << src[r]) - 1u;
        for (uint32_t k = 0; k < wm[r]; k++) {
            if (k < sn[r]) {
                const uint2 v = pm[t0[r] + k];
                const uint32_t dst = v.x + (uint32_t)__popc(v.y & lower);
                p.vals[dst] = id;
            }
        }
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < STRIP_ROUNDS; r++)
        for (uint32_t k = 0; k < wm[r]; k++) {
            if (k < sn[r]) {
                const uint2 v = pm[t0[r] + k];
                // the highest owner of the tile's group: one update per (batch, tile)
                if ((v.y >> src[r]) == 1u) pm[t0[r] + k] = make_uint2(v.x + (uint32_t)__popc(v.y), 0u);
            }
        }
    __syncwarp();
}

#ifdef D3GA_SCATTER_TRACE
// development trace: per (row, warp) start ns, set-up done ns, end ns, SM id | batches << 32
__device__ unsigned long long g_sc_trace[8192 * SC_WARPS * 4];
__device__ __forceinline__ unsigned long long sc_global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

__global__ void __launch_bounds__(SC_WARPS * 32) direct_scatter_kernel(const DirectParams p, int gy, int band_rows) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ ScatterWarp s_warp[SC_WARPS];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, row = blockIdx.x;
#ifdef D3GA_SCATTER_TRACE
    const unsigned long long tr_t0 = sc_global_ns();
    unsigned long long tr_t1 = 0;
    uint32_t tr_batches = 0;
#endif
    if (p.hdr->overflow) return;
    // this CTA's band of tile rows [yb0, yb1) (the whole image when it fits the shared memory budget); tiles, tile rows
    // and blocks below are relative to the band
    const uint32_t gx = (uint32_t)p.gx, yb0 = blockIdx.y * (uint32_t)band_rows, yb1 = min((uint32_t)gy, yb0 + (uint32_t)band_rows);
    const uint32_t tbase = yb0 * gx, ntiles = (yb1 - yb0) * gx;
    uint2 *pm = reinterpret_cast<uint2 *>(smem);   // [tiles of the band] x: next free slot of the tile's run, y: owners of the batch on it
    {
        // (tbase is a multiple of 4: bands start on a multiple of 16 tile rows; reads past the band's last tile stay
        // inside the rows' tl_pad)
        const uint4 *srow = reinterpret_cast<const uint4 *>(p.seg + (size_t)(row / p.rows_per_seg) * p.tl_pad + tbase);
        const uint2 *mrow = reinterpret_cast<const uint2 *>(p.before + (size_t)row * p.tl_pad + tbase);
        for (uint32_t i = tid; i < (ntiles + 3) / 4; i += SC_WARPS * 32) {
            const uint4 v = __ldg(srow + i);
            const uint2 m = __ldg(mrow + i);
            uint4 *o = reinterpret_cast<uint4 *>(pm + 4 * i);
            o[0] = make_uint4(v.x + (m.x & 0xffffu), 0u, v.y + (m.x >> 16), 0u);
            o[1] = make_uint4(v.z + (m.y & 0xffffu), 0u, v.w + (m.y >> 16), 0u);
        }
    }
    __syncthreads();
#ifdef D3GA_SCATTER_TRACE
    tr_t1 = sc_global_ns();
#endif
    if (warp << BLK_SHIFT >= yb1 - yb0) return;   // fewer blocks of tile rows than warps
    ScatterWarp &sw = s_warp[warp];
    const uint32_t slot0 = row * p.chunk, nb = p.chunk / 32, below = (1u << lane) - 1u;
    auto load = [&](uint32_t b) {
        const uint32_t s = slot0 + b * 32 + lane;
        return (b < nb && s < (uint32_t)p.P) ? p.rec[s] : make_uint4(0, 0, 0, 0);
    };
    uint32_t qh = 0, qn = 0;   // the FIFO's head and fill (warp-uniform)
    // records: two batches ahead in registers (an L1 prefetch further ahead was measured: no gain)
    uint4 rec_a = load(0), rec_b = load(1), rec = make_uint4(0, 0, 0, 0);
    uint32_t b = 0;
    uint32_t ya = 0, yb = 0, first = 0, last = 0, c = 0;   // this lane's rect inside the band: tile rows, this warp's blocks
    uint32_t kcur = 1, kend = 0;   // blocks of the current batch still to be queued (tall rects only)
    auto enqueue = [&](uint32_t mask, bool mine, uint32_t blk) {   // one entry per lane of `mask`, in lane order
        if (mine) {
            const uint32_t y0e = max(ya, blk << BLK_SHIFT), y1e = min(yb, (blk + 1) << BLK_SHIFT);
            sw.queue[(qh + qn + (uint32_t)__popc(mask & below)) & (QUEUE - 1)] =
                make_uint4((rec.x & 0xffffu) | (y0e << 16), (rec.y & 0xffffu) | ((y1e - y0e) << 16), rec.z, rec.w);
        }
        qn += (uint32_t)__popc(mask);
        __syncwarp();
    };
    for (;;) {
        while (qn < 32) {
            if (kcur <= kend) {
                // rects that span several of this warp's blocks: block by block.  Entries that can share a tile lie in
                // the same block, so inside every block they are queued in lane order = depth order.
                const bool mine = c && first <= kcur && kcur <= last;
                const uint32_t m = __ballot_sync(0xffffffffu, mine);
                if (m) enqueue(m, mine, kcur);
                kcur += SC_WARPS;
                continue;
            }
            if (b >= nb) break;
            rec = rec_a;
            rec_a = rec_b;
            rec_b = load(b + 2);
            b++;
            // the lane's rect clipped to the band, the blocks of tile rows it spans, and those of them that are this warp's
            const uint32_t w = rec.y & 0xffffu, y0 = rec.x >> 16, y1 = y0 + (rec.y >> 16);
            ya = max(y0, yb0); yb = min(y1, yb1);
            c = 0;
            if (w && yb > ya) {
                ya -= yb0; yb -= yb0;
                const uint32_t blk0 = ya >> BLK_SHIFT, blk1 = (yb - 1) >> BLK_SHIFT;
                first = blk0 + ((warp - blk0) & (SC_WARPS - 1));
                if (first <= blk1) { c = (blk1 - first) / SC_WARPS + 1; last = first + (c - 1) * SC_WARPS; }
            }
            const uint32_t any = __ballot_sync(0xffffffffu, c > 0);
            if (!any) continue;
            if (__ballot_sync(0xffffffffu, c > 1)) {
                kcur = __reduce_min_sync(0xffffffffu, c ? first : 0xffffffffu);
                kend = __reduce_max_sync(0xffffffffu, c ? last : 0u);
                continue;
            }
            // at most one entry per lane (rects under 13 tile rows: practically always at 1024^2)
            enqueue(any, c != 0, first);
        }
        if (qn == 0) break;
        const uint32_t n = min(qn, 32u);
        scatter_batch(p, sw, pm, qh, n, lane);
        qh += n;
        qn -= n;
#ifdef D3GA_SCATTER_TRACE
        tr_batches++;
#endif
    }
#ifdef D3GA_SCATTER_TRACE
    if (lane == 0 && row < 8192 && blockIdx.y == 0) {
        uint32_t smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        unsigned long long *o = g_sc_trace + 4 * ((size_t)row * SC_WARPS + warp);
        o[0] = tr_t0; o[1] = tr_t1; o[2] = sc_global_ns(); o[3] = ((unsigned long long)tr_batches << 32) | smid;
    }
#endif
}

#ifdef D3GA_SCATTER_TRACE
}  // namespace
}  // namespace d3ga
extern "C" int d3ga_debug_scatter_trace(unsigned long long *out, int rows) {
    return cudaMemcpyFromSymbol(out, d3ga::g_sc_trace, sizeof(unsigned long long) * 4 * d3ga::SC_WARPS * (size_t)rows) == cudaSuccess ? 0 : -2;
}
namespace d3ga {
namespace {
#endif

// Column sums of M over segments of rows (two tiles per thread: M's 16-bit counts are read as pairs); the last CTA of a
// column block then scans its columns over the segments (exclusive, in place) and writes the per-tile totals.
__global__ void __launch_bounds__(COL_THREADS) direct_segsum_kernel(const uint32_t *__restrict__ M32, uint2 *seg, uint2 *tot,
                                                                    uint32_t *tickets, uint32_t rows, uint32_t npairs,
                                                                    uint32_t nseg, uint32_t rows_per_seg) {
    __shared__ uint32_t s_last;
    const uint32_t pair = blockIdx.x * COL_THREADS + threadIdx.x, s = blockIdx.y;
    const uint32_t r0 = s * rows_per_seg, r1 = min(rows, r0 + rows_per_seg);
    if (pair < npairs) {
        uint32_t lo = 0, hi = 0;
        for (uint32_t r = r0; r < r1; r += 8) {
            uint32_t m[8];
#pragma unroll
            for (int u = 0; u < 8; u++) m[u] = r + u < r1 ? __ldcg(M32 + (size_t)(r + u) * npairs + pair) : 0u;
#pragma unroll
            for (int u = 0; u < 8; u++) { lo += m[u] & 0xffffu; hi += m[u] >> 16; }
        }
        seg[(size_t)s * npairs + pair] = make_uint2(lo, hi);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&tickets[blockIdx.x], 1u) == nseg - 1;
    __syncthreads();
    if (!s_last || pair >= npairs) return;
    __threadfence();
    uint2 run = make_uint2(0, 0);
    for (uint32_t q0 = 0; q0 < nseg; q0 += 8) {
        uint2 v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = q0 + u < nseg ? __ldcg(seg + (size_t)(q0 + u) * npairs + pair) : make_uint2(0, 0);
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (q0 + u < nseg) seg[(size_t)(q0 + u) * npairs + pair] = run;
            run.x += v[u].x; run.y += v[u].y;
        }
    }
    tot[pair] = run;
}

// One CTA: start of every tile's run in the sorted list (exclusive scan of the totals in tile order), ranges, the plan
// words of the header and the forward blend's tile order (longest list first).
__global__ void __launch_bounds__(256) direct_plan_kernel(RasterHeader *hdr, const uint32_t *__restrict__ tot, uint32_t *start,
                                                          int num_tiles, uint2 *ranges, uint32_t *order_fwd, int sort_bits,
                                                          int num_passes, int grid_x) {
    extern __shared__ uint32_t s_cnt[];   // [num_tiles]
    __shared__ uint32_t s_tmp[8];
    __shared__ uint32_t s_bins[257];
    const int tid = threadIdx.x;
    for (int i0 = 0; i0 < num_tiles; i0 += 256 * 8) {
        uint32_t v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int i = i0 + u * 256 + tid; v[u] = i < num_tiles ? __ldcg(tot + i) : 0u; }
#pragma unroll
        for (int u = 0; u < 8; u++) { const int i = i0 + u * 256 + tid; if (i < num_tiles) s_cnt[i] = v[u]; }
    }
    __syncthreads();
    // Two CTAs, side by side: CTA 0 scans the totals (what the next kernels wait for), CTA 1 orders the tiles for the
    // forward blend (needed much later).  One after the other in a single CTA they were 20 us of serial latency.
    if (blockIdx.x == 1) {
        // Launch order of the forward blend: longest WALK first.  A tile's list is walked until its last pixel saturates:
        // inside the body that happens after ~1.5 k entries however long the list is, while a tile on the silhouette (a
        // neighbour with a much shorter list: some of its pixels never saturate) is walked to the end.  Ordering by list
        // length alone started those tiles ~640 slots late and their 8 serial warps ended the kernel alone (a per-warp
        // %globaltimer trace: 126 us of chain starting at t = 40 us); the estimate below starts them first.
        const uint32_t *c = s_cnt;
        const int gx = grid_x, gy = num_tiles / grid_x;
        auto walk = [c, gx, gy](int t) {
            const uint32_t n = c[t];
            constexpr uint32_t SAT = 1536u;
            if (n <= SAT) return n;
            const int x = t % gx, y = t / gx;
            uint32_t mn = n;
            if (x > 0) mn = min(mn, c[t - 1]);
            if (x + 1 < gx) mn = min(mn, c[t + 1]);
            if (y > 0) mn = min(mn, c[t - gx]);
            if (y + 1 < gy) mn = min(mn, c[t + gx]);
            return 4u * mn < n ? n : SAT + (n - SAT) / 8u;
        };
        order_tiles_descending(num_tiles, walk, order_fwd, s_bins);
        return;
    }
    const int per = (num_tiles + 255) / 256, t0 = min(num_tiles, tid * per), t1 = min(num_tiles, t0 + per);
    uint32_t local = 0;
    for (int t = t0; t < t1; t++) local += s_cnt[t];
    // exclusive scan of one value per thread over the CTA's 256 threads
    const uint32_t lane = tid & 31, warp = tid >> 5;
    uint32_t incl = local;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    if (lane == 31) s_tmp[warp] = incl;
    __syncthreads();
    uint32_t run = incl - local;
    for (uint32_t w = 0; w < warp; w++) run += s_tmp[w];
    for (int t = t0; t < t1; t++) {
        const uint32_t c = s_cnt[t];
        start[t] = run;
        ranges[t] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
        run += c;
    }
    if (tid == 0) {
        hdr->final_sel = 0u;
        hdr->sort_bits = (uint32_t)sort_bits;
        hdr->num_passes = (uint32_t)num_passes;
    }
    if (tid < SORT_MAX_PASSES) { hdr->pass_skip[tid] = 1u; hdr->pass_src[tid] = 0u; }   // no radix pass runs
}

// Segment offsets += start[tile] (they become absolute slots); before[row][tile] = exclusive prefix of M over the rows of
// the row's segment (at most rows_per_seg * chunk <= 65535 duplicates: 16 bits).
__global__ void __launch_bounds__(COL_THREADS) direct_base_kernel(const uint32_t *__restrict__ M32, uint32_t *before32, uint2 *seg,
                                                                  const uint2 *__restrict__ start, uint32_t rows,
                                                                  uint32_t npairs, uint32_t rows_per_seg) {
    const uint32_t pair = blockIdx.x * COL_THREADS + threadIdx.x, s = blockIdx.y;
    const uint32_t r0 = s * rows_per_seg, r1 = min(rows, r0 + rows_per_seg);
    if (pair >= npairs || r0 >= r1) return;
    {
        uint2 so = seg[(size_t)s * npairs + pair];
        const uint2 st = start[pair];
        so.x += st.x; so.y += st.y;
        seg[(size_t)s * npairs + pair] = so;
    }
    uint32_t lo = 0, hi = 0;
    for (uint32_t r = r0; r < r1; r += 8) {
        uint32_t m[8];
#pragma unroll
        for (int u = 0; u < 8; u++) m[u] = r + u < r1 ? __ldcg(M32 + (size_t)(r + u) * npairs + pair) : 0u;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (r + u < r1) before32[(size_t)(r + u) * npairs + pair] = lo | (hi << 16);
            lo += m[u] & 0xffffu; hi += m[u] >> 16;
        }
    }
}

std::atomic<int> g_views_in_flight{1};

uint32_t chunk_override() {
    static const uint32_t v = [] {
        const char *e = getenv("D3GA_DIRECT_CHUNK");
        long c = e ? atol(e) : 0;
        if (c <= 0) return 0u;
        if (c > DIRECT_MAX_CHUNK) c = DIRECT_MAX_CHUNK;
        return (uint32_t)((c + 31) / 32 * 32);
    }();
    return v;
}
DirectParams make_params(int P, int gx, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                         uint64_t capacity, const DirectPlan &pl) {
    DirectParams p;
    p.hdr = reinterpret_cast<RasterHeader *>(geom + GL.header);
    p.P = P; p.gx = gx; p.P_pad = (uint32_t)GL.p_pad; p.tl_pad = pl.tl_pad; p.chunk = pl.chunk; p.rows = pl.rows;
    p.rows_per_seg = pl.rows_per_seg;
    p.capacity = capacity;
    p.dvals = reinterpret_cast<const uint32_t *>(geom + GL.dvals);
    p.grec = reinterpret_cast<const uint4 *>(geom + GL.grec);
    p.rec = reinterpret_cast<uint4 *>(geom + GL.ordered);
    p.M = reinterpret_cast<uint16_t *>(binning + BL.vals1);
    p.seg = reinterpret_cast<const uint32_t *>(binning + BL.keys1 + pl.seg_off);
    p.before = reinterpret_cast<const uint16_t *>(binning + BL.keys1 + pl.before_off);
    p.vals = reinterpret_cast<uint32_t *>(binning + BL.vals0);
    return p;
}

int set_smem_attrs() {
    static std::atomic<uint64_t> devices{0};
    return once_per_device(devices, []() -> int {
        const int scatter = 200 * 1024, plan = 4 * DIRECT_MAX_TILES, count = 4 * DIRECT_MAX_CELLS;
        D3GA_CUDA(cudaFuncSetAttribute(direct_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, count));
        D3GA_CUDA(cudaFuncSetAttribute(direct_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, scatter));
        D3GA_CUDA(cudaFuncSetAttribute(direct_plan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, plan));
        return D3GA_OK;
    });
}

// Stage 1 decides the plan -- it depends on the views-in-flight hint, a process-wide setting another thread may change at
// any moment -- and files it under the geometry workspace it ran on; stage 2 (and d3ga_raster_sorted_keys) take the plan
// from there instead of deciding again, so the two stages of a view can never disagree about rows, segments or offsets.
struct PlanRecord {
    int dev;
    const void *geom;
    int P, gx, gy;
    uint64_t capacity;
    DirectPlan plan;
    const void *packed_colors;
};
constexpr int PLAN_RECORDS = 512;   // workspaces between their stage 1 and their last use of the plan
std::mutex g_plan_mu;
PlanRecord g_plans[PLAN_RECORDS];
unsigned g_plan_next = 0;

}  // namespace

DirectPlan direct_plan(int P, int gx, int gy, uint64_t capacity) {
    DirectPlan pl = {};
    const uint64_t tl = (uint64_t)gx * gy, cap = capacity < 1 ? 1 : capacity;
    if (sort_mode() != SORT_DIRECT || P < 1 || tl < 1 || tl > DIRECT_MAX_TILES || gx > 2047 ||
        (uint64_t)(gx + 1) * (gy + 1) > DIRECT_MAX_CELLS)
        return pl;
    pl.tl_pad = (uint32_t)((tl + 7) & ~7ull);
    pl.colblocks = (pl.tl_pad / 2 + COL_THREADS - 1) / COL_THREADS;
    // The scatter keeps 8 B per tile in shared memory.  Images whose tiles do not fit a ~36 KB budget (six CTAs per SM)
    // are cut into bands of tile rows (a multiple of 16: four blocks per warp-round), one CTA per (row, band).
    pl.band_rows = (uint32_t)gy;
    pl.bands = 1;
    if (8ull * pl.tl_pad > SCATTER_SMEM_BUDGET) {
        uint64_t br = SCATTER_SMEM_BUDGET / (8ull * gx) / 16 * 16;
        if (br < 16) br = 16;
        if (8ull * br * gx + 64 > 200 * 1024) return pl;   // a single band of 16 tile rows does not fit: not this path
        pl.band_rows = (uint32_t)br;
        pl.bands = (uint32_t)((gy + br - 1) / br);
    }
    pl.band_tiles_pad = (uint32_t)((((uint64_t)(pl.bands > 1 ? pl.band_rows : gy) * gx) + 7) & ~7ull);
    // CTAs that are resident at once (+1 KB the hardware reserves per CTA, 228 KB and at most 16 of these CTAs per SM);
    // the row length is what fills them once
    uint64_t per_sm = (228ull * 1024) / (8ull * pl.band_tiles_pad + 1024 + sizeof(ScatterWarp) * SC_WARPS);
    if (per_sm > 16) per_sm = 16;
    if (per_sm < 1) return pl;
    uint64_t chunk = chunk_override();
    if (!chunk) {
        // ... once for a single view; a caller that keeps n views in flight on separate streams (d3ga_set_views_in_flight)
        // gets rows n times as long: each view's scatter then holds about 1/n of the shared memory and warp slots for n
        // times as long instead of all of them, which leaves the other views' kernels room beside it
        int vif = g_views_in_flight.load(std::memory_order_relaxed);
        if (vif < 1) vif = 1;
        const uint64_t sms = (uint64_t)device_sm_count();
        uint64_t rows_target = sms * per_sm / pl.bands / (uint64_t)vif;
        if (rows_target < sms) rows_target = sms;
        chunk = ((uint64_t)P + rows_target - 1) / rows_target;
        chunk = (chunk + 31) / 32 * 32;
        if (chunk < 256) chunk = 256;
        if (chunk > DIRECT_MAX_CHUNK) chunk = DIRECT_MAX_CHUNK;
    }
    // scratch: M (2 B x rows x tiles) in the unused value buffer (4 B x capacity); segment offsets, totals, starts and
    // the column-block tickets in the unused key buffer (8 B x capacity).  Longer rows when M would not fit.
    const uint64_t rows_max = 2 * cap / pl.tl_pad;
    if (rows_max < 1) return pl;
    if (((uint64_t)P + chunk - 1) / chunk > rows_max) chunk = ((((uint64_t)P + rows_max - 1) / rows_max) + 31) / 32 * 32;
    if (chunk > DIRECT_MAX_CHUNK) return pl;
    pl.chunk = (uint32_t)chunk;
    pl.rows = (uint32_t)(((uint64_t)P + chunk - 1) / chunk);
    // a segment's duplicates of one tile must fit 16 bits: rows_per_seg * chunk <= 65535
    const uint32_t rps_max = 65535u / pl.chunk;
    uint32_t nseg = pl.rows < (uint32_t)DIRECT_MAX_SEGS ? pl.rows : (uint32_t)DIRECT_MAX_SEGS;
    pl.rows_per_seg = (pl.rows + nseg - 1) / nseg;
    if (pl.rows_per_seg > rps_max) pl.rows_per_seg = rps_max;
    pl.nseg = (pl.rows + pl.rows_per_seg - 1) / pl.rows_per_seg;
    pl.seg_off = 0;
    pl.tot_off = align256(pl.seg_off + 4ull * pl.nseg * pl.tl_pad);
    pl.start_off = align256(pl.tot_off + 4ull * pl.tl_pad);
    pl.ticket_off = align256(pl.start_off + 4ull * pl.tl_pad);
    pl.before_off = align256(pl.ticket_off + 4ull * pl.colblocks);
    if (pl.before_off + 2ull * pl.rows * pl.tl_pad > 8ull * cap) return pl;
    pl.ok = 1;
    return pl;
}

// The sorted (tile | depth) key list is not needed by the blend kernels (they read the Gaussian ids and the ranges), and
// storing 8 scattered bytes per duplicate costs more than the rest of the scatter, so the direct path does not write it.
// When it is asked for (tests, inspection) it is rebuilt from what defines it: the tile of slot i comes from the ranges,
// the depth bits from the Gaussian in that slot.
__global__ void __launch_bounds__(256) sorted_keys_kernel(const RasterHeader *hdr, const uint2 *__restrict__ ranges,
                                                          const uint32_t *__restrict__ vals, const float *__restrict__ depth,
                                                          uint64_t *keys) {
    if (hdr->overflow) return;
    const uint32_t tile = blockIdx.x;
    const uint2 r = ranges[tile];
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += 256) keys[i] = ((uint64_t)tile << 32) | __float_as_uint(depth[vals[i]]);
}

int launch_sorted_keys(int num_tiles, const uint8_t *geom, const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL,
                       const uint8_t *image, const ImgLayout &IL, uint64_t *keys_out, cudaStream_t stream) {
    if (num_tiles < 1) return D3GA_OK;
    sorted_keys_kernel<<<num_tiles, 256, 0, stream>>>(
        reinterpret_cast<const RasterHeader *>(geom + GL.header), reinterpret_cast<const uint2 *>(image + IL.ranges),
        reinterpret_cast<const uint32_t *>(binning + BL.vals0), reinterpret_cast<const float *>(geom + GL.depth), keys_out);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

// The radix paths sort real keys: copy them out of whichever buffer the last pass wrote (the header knows).
__global__ void __launch_bounds__(256) copy_sorted_keys_kernel(const RasterHeader *hdr, const uint64_t *__restrict__ keys0,
                                                               const uint64_t *__restrict__ keys1, uint64_t *out) {
    if (hdr->overflow) return;
    const uint64_t *src = hdr->final_sel ? keys1 : keys0;
    const uint64_t D = hdr->num_rendered, stride = (uint64_t)gridDim.x * 256;
    for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < D; i += stride) out[i] = src[i];
}

int launch_copy_sorted_keys(const uint8_t *geom, const GeomLayout &GL, const uint8_t *binning, const BinLayout &BL,
                            uint64_t capacity, uint64_t *keys_out, cudaStream_t stream) {
    uint64_t grid = (capacity + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    if (grid < 1) grid = 1;
    copy_sorted_keys_kernel<<<(unsigned)grid, 256, 0, stream>>>(
        reinterpret_cast<const RasterHeader *>(geom + GL.header), reinterpret_cast<const uint64_t *>(binning + BL.keys0),
        reinterpret_cast<const uint64_t *>(binning + BL.keys1), keys_out);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

void remember_direct_plan(const void *geom, int P, int gx, int gy, uint64_t capacity, const DirectPlan &pl,
                          const void *packed_colors) {
    int dev = 0;
    (void)cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_plan_mu);
    for (int i = 0; i < PLAN_RECORDS; i++)
        if (g_plans[i].geom == geom && g_plans[i].dev == dev) {
            g_plans[i] = PlanRecord{dev, geom, P, gx, gy, capacity, pl, packed_colors};
            return;
        }
    g_plans[g_plan_next++ % PLAN_RECORDS] = PlanRecord{dev, geom, P, gx, gy, capacity, pl, packed_colors};
}
bool recall_direct_plan(const void *geom, int P, int gx, int gy, uint64_t capacity, DirectPlan *pl) {
    int dev = 0;
    (void)cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_plan_mu);
    for (int i = 0; i < PLAN_RECORDS; i++) {
        const PlanRecord &r = g_plans[i];
        if (r.geom == geom && r.dev == dev) {
            if (r.P != P || r.gx != gx || r.gy != gy || r.capacity != capacity) return false;
            *pl = r.plan;
            return true;
        }
    }
    return false;
}

const void *recall_packed_colors(const void *geom) {
    int dev = 0;
    (void)cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_plan_mu);
    for (int i = 0; i < PLAN_RECORDS; i++)
        if (g_plans[i].geom == geom && g_plans[i].dev == dev) return g_plans[i].packed_colors;
    return nullptr;
}

void set_views_in_flight(int n) { g_views_in_flight.store(n < 1 ? 1 : (n > 64 ? 64 : n), std::memory_order_relaxed); }

// Stage 1 tail: the count rows, the depth-ordered records and the duplicate total.
int launch_direct_count(int P, int gx, int gy, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                        uint64_t capacity, const DirectPlan &pl, cudaStream_t stream) {
    const DirectParams p = make_params(P, gx, geom, GL, binning, BL, capacity, pl);
    int rc = set_smem_attrs();
    if (rc) return rc;
    ProfScope prof(K_DIRECT_COUNT, stream);
    direct_count_kernel<<<pl.rows, COUNT_WARPS * 32, 4 * (size_t)(gx + 1) * (gy + 1), stream>>>(p, gy, gx * gy);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

// Stage 2: offsets from the count rows, then the scatter.
int launch_direct_scatter(int P, int gx, int gy, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                          uint64_t capacity, int sort_bits, uint8_t *image, const ImgLayout &IL, const DirectPlan &pl,
                          cudaStream_t stream) {
    const DirectParams p = make_params(P, gx, geom, GL, binning, BL, capacity, pl);
    const int num_tiles = gx * gy;
    uint8_t *scratch = binning + BL.keys1;
    const uint32_t *M32 = reinterpret_cast<const uint32_t *>(p.M);
    uint2 *seg = reinterpret_cast<uint2 *>(scratch + pl.seg_off);
    uint2 *tot = reinterpret_cast<uint2 *>(scratch + pl.tot_off);
    uint2 *start = reinterpret_cast<uint2 *>(scratch + pl.start_off);
    uint32_t *tickets = reinterpret_cast<uint32_t *>(scratch + pl.ticket_off);
    const uint32_t npairs = pl.tl_pad / 2;
    int rc = set_smem_attrs();
    if (rc) return rc;
    D3GA_CUDA(cudaMemsetAsync(tickets, 0, 4ull * pl.colblocks, stream));
    const dim3 colgrid(pl.colblocks, pl.nseg);
    {
        ProfScope prof(K_DIRECT_SEGSUM, stream);
        direct_segsum_kernel<<<colgrid, COL_THREADS, 0, stream>>>(M32, seg, tot, tickets, pl.rows, npairs, pl.nseg,
                                                                  pl.rows_per_seg);
    }
    D3GA_LAUNCH_CHECK();
    {
        ProfScope prof(K_DIRECT_PLAN, stream);
        direct_plan_kernel<<<2, 256, 4 * (size_t)num_tiles, stream>>>(
            p.hdr, reinterpret_cast<const uint32_t *>(tot), reinterpret_cast<uint32_t *>(start), num_tiles,
            reinterpret_cast<uint2 *>(image + IL.ranges), reinterpret_cast<uint32_t *>(image + IL.order_fwd), sort_bits,
            (sort_bits + 7) / 8, gx);
    }
    D3GA_LAUNCH_CHECK();
    {
        ProfScope prof(K_DIRECT_BASE, stream);
        direct_base_kernel<<<colgrid, COL_THREADS, 0, stream>>>(M32, reinterpret_cast<uint32_t *>(scratch + pl.before_off), seg,
                                                                start, pl.rows, npairs, pl.rows_per_seg);
    }
    D3GA_LAUNCH_CHECK();
    {
        ProfScope prof(K_DIRECT_SCATTER, stream);
        direct_scatter_kernel<<<dim3(pl.rows, pl.bands), 32 * SC_WARPS, 8 * (size_t)pl.band_tiles_pad, stream>>>(p, gy, (int)pl.band_rows);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
