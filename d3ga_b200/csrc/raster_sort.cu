// raster_sort.cu -- rows R4 (radix sort of the (tile|depth) keys) and R5 (per-tile ranges) of SURVEY.md section 8.
//
// Replaces cub::DeviceRadixSort::SortPairs(keys, values, num_rendered, 0, 32 + bit) and identifyTileRanges inside
// diff_gaussian_rasterization._C.rasterize_gaussians (reached from renderer.py:130-141).
//
// Design: a onesweep least-significant-digit radix sort, 8-bit digits, stable (equal keys keep Gaussian order, as
// the upstream sort does), with everything that depends on the duplicate count D read from DEVICE memory so the
// host never waits for it:
//   1. one histogram kernel counts all passes' digits in a single read of the keys; its last CTA turns the counts
//      into per-pass exclusive bin offsets and a *plan*: a pass in which every key has the same digit (constant
//      float exponent byte of the depth, unused tile bits) is skipped -- the output is bit-identical to a full sort;
//   2. one kernel per remaining pass: each CTA takes a 4096-pair tile (dynamic ticket), ranks its keys with
//      warp-level match-any multi-split, publishes its digit counts and resolves its global offsets with a
//      decoupled look-back (one word per digit per tile), reorders the tile in shared memory and writes runs of
//      equal digits with coalesced stores.  HBM traffic per pass: 12 B read + 12 B written per pair.
//   3. ranges: neighbour compare of key >> 32 on the sorted list.
// Algorithmic bytes: 8 D (histogram) + 24 D per executed pass + 8 D + 8 Tl (ranges).
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace d3ga {

namespace {

// look-back word: bit 31 set = inclusive prefix in the low 31 bits; else bit 30 set = tile aggregate (<= SORT_TILE) in
// the low 30 bits; 0 = not published yet
constexpr uint32_t ST_INC = 1u << 31, ST_AGG = 1u << 30;
__device__ __forceinline__ bool st_ready(uint32_t s) { return (s & (ST_INC | ST_AGG)) != 0u; }
__device__ __forceinline__ bool st_inclusive(uint32_t s) { return (s & ST_INC) != 0u; }
__device__ __forceinline__ uint32_t st_value(uint32_t s) { return (s & ST_INC) ? (s & ~ST_INC) : (s & ~ST_AGG); }
constexpr int HIST_THREADS = 256;
constexpr int HIST_ITEMS = 8;

__global__ void __launch_bounds__(HIST_THREADS) sort_histogram_kernel(RasterHeader *hdr, const uint64_t *keys,
                                                                      uint32_t *ghist, uint32_t *bins, int sort_bits,
                                                                      int num_passes, int first_pass) {
    __shared__ uint32_t s_hist[SORT_MAX_PASSES][256];
    __shared__ uint32_t s_tmp[8];
    __shared__ uint32_t s_flag;
    if (hdr->overflow) return;
    const uint32_t D = hdr->num_rendered;
    const int tid = threadIdx.x;
    for (int p = 0; p < num_passes; p++) s_hist[p][tid] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * HIST_THREADS * HIST_ITEMS;
    for (uint64_t base = (uint64_t)blockIdx.x * HIST_THREADS * HIST_ITEMS; base < D; base += stride) {
        uint64_t k[HIST_ITEMS];
        bool ok[HIST_ITEMS];
#pragma unroll
        for (int i = 0; i < HIST_ITEMS; i++) {
            const uint64_t idx = base + (uint64_t)i * HIST_THREADS + tid;
            ok[i] = idx < D;
            k[i] = ok[i] ? __ldcs(keys + idx) : 0;
        }
#pragma unroll
        for (int i = 0; i < HIST_ITEMS; i++) {
            const uint32_t active = __ballot_sync(0xffffffffu, ok[i]);
            if (!active) continue;
            const uint32_t lane = lane_id();
            // Depth digits (low 32 bits): consecutive keys of one Gaussian share the depth, so only the first lane of
            // each run counts, with the run length -- a fifth of the shared-memory atomics on typical scenes.
            const uint32_t depth = (uint32_t)k[i];
            const uint32_t prev = __shfl_up_sync(0xffffffffu, depth, 1);
            const bool head = ok[i] && (lane == 0 || prev != depth);
            const uint32_t heads = __ballot_sync(0xffffffffu, head);
            const uint32_t above = lane == 31 ? 0u : (heads & ~((2u << lane) - 1u));
            const uint32_t run = (above ? (uint32_t)(__ffs(above) - 1) : (uint32_t)__popc(active)) - lane;
            for (int p = first_pass; p < num_passes; p++) {
                const int shift = 8 * p;
                if (shift < 32) {
                    if (head) atomicAdd(&s_hist[p][(depth >> shift) & 0xffu], run);
                    continue;
                }
                // tile digits: neighbouring lanes hit neighbouring bins; a whole warp in one bin adds once
                const uint32_t nb = min(8, sort_bits - shift);
                const uint32_t d = (uint32_t)(k[i] >> shift) & ((1u << nb) - 1u);
                const uint32_t d0 = __shfl_sync(0xffffffffu, d, __ffs(active) - 1);
                const bool uniform = __all_sync(0xffffffffu, !ok[i] || d == d0);
                if (uniform) {
                    if (lane == (uint32_t)(__ffs(active) - 1)) atomicAdd(&s_hist[p][d0], __popc(active));
                } else if (ok[i]) {
                    atomicAdd(&s_hist[p][d], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int p = 0; p < num_passes; p++) {
        const uint32_t c = s_hist[p][tid];
        if (c) atomicAdd(&ghist[p * 256 + tid], c);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) s_flag = atomicAdd(&hdr->hist_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_flag) return;
    // ---- last CTA: bin offsets and the pass plan ---------------------------------------------------------------------
    __threadfence();
    uint32_t src = 0;
    for (int p = 0; p < num_passes; p++) {
        const uint32_t c = __ldcg(ghist + p * 256 + tid);
        // passes below first_pass (the depth digits) are left to the per-tile shared-memory sort
        const int all_in_one = __syncthreads_or(D == 0 || c == D || p < first_pass);
        const uint32_t ex = block_excl_scan_256(c, s_tmp, nullptr);
        bins[p * 256 + tid] = ex;
        if (tid == 0) {
            hdr->pass_skip[p] = all_in_one ? 1u : 0u;
            hdr->pass_src[p] = src;
        }
        if (!all_in_one) src ^= 1u;
    }
    if (tid == 0) {
        hdr->final_sel = src;
        hdr->sort_bits = (uint32_t)sort_bits;
        hdr->num_passes = (uint32_t)num_passes;
    }
}

#ifndef LOOKBACK_WIDTH
#define LOOKBACK_WIDTH 8   // predecessors read per round trip of the look-back
#endif
constexpr int SORT_WARPS = SORT_THREADS / 32;
template <typename KeyT>
constexpr int sort_smem_bytes() {
    return SORT_TILE * (int)sizeof(KeyT) + SORT_TILE * 4 + 4 * (SORT_WARPS * 256 + 256 + 256 + 16);
}

// PLAN 0: the (tile|depth) duplicate list, count and plan from the header (num_rendered, pass_*).
// PLAN 1: the 32-bit depth keys of the Gaussians (depth-first binning), count = n_fixed, plan in dep_*.
#ifdef D3GA_SORT_TRACE
// development trace of the depth-sort passes: per (pass, tile) ns at start / keys loaded / ranked / look-back done / end
__device__ unsigned long long g_sort_trace[4 * 512 * 8];
__device__ __forceinline__ unsigned long long sort_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define SORT_TRACE_MARK(i) if (PLAN == 1 && threadIdx.x == 0 && tr_tile < 512) g_sort_trace[((size_t)pass * 512 + tr_tile) * 8 + (i)] = sort_ns();
#else
#define SORT_TRACE_MARK(i)
#endif

template <typename KeyT, int PLAN>
__global__ void __launch_bounds__(SORT_THREADS) sort_pass_kernel(RasterHeader *hdr, KeyT *keys0, KeyT *keys1,
                                                                 uint32_t *vals0, uint32_t *vals1,
                                                                 const uint32_t *bins, uint32_t *tile_state,
                                                                 uint32_t tiles_cap, int pass, int shift,
                                                                 uint32_t mask, uint32_t n_fixed) {
    extern __shared__ __align__(16) uint8_t smem[];
    KeyT *s_keys = reinterpret_cast<KeyT *>(smem);
    uint32_t *s_vals = reinterpret_cast<uint32_t *>(smem + SORT_TILE * sizeof(KeyT));
    uint32_t *s_whist = s_vals + SORT_TILE;         // [warps][256] per-warp digit counts -> exclusive over warps
    uint32_t *s_start = s_whist + SORT_WARPS * 256; // [256] first slot of each digit inside the tile
    uint32_t *s_gbase = s_start + 256;              // [256] global slot of s_start[d] minus s_start[d]
    uint32_t *s_misc = s_gbase + 256;               // [0..7] scan scratch, [8] tile id

#ifdef D3GA_SORT_TRACE
    uint32_t tr_tile = blockIdx.x;
    SORT_TRACE_MARK(0)
#endif
    if (PLAN == 0 ? (hdr->overflow || hdr->pass_skip[pass]) : hdr->dep_skip[pass]) return;
    const uint32_t D = PLAN == 0 ? hdr->num_rendered : n_fixed;
    uint32_t *ticket = PLAN == 0 ? &hdr->pass_ticket[pass] : &hdr->dep_ticket[pass];
    const uint32_t src = PLAN == 0 ? hdr->pass_src[pass] : hdr->dep_src[pass];
    const int tid = threadIdx.x;
    const uint32_t lane = lane_id(), warp = tid >> 5;
    if (tid == 0) s_misc[8] = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_misc[8];
    const uint64_t tile_base = (uint64_t)tile * SORT_TILE;
    if (tile_base >= D) return;
#ifdef D3GA_SORT_TRACE
    tr_tile = tile;
    SORT_TRACE_MARK(1)
#endif
    const uint32_t n_valid = (uint32_t)min((uint64_t)SORT_TILE, (uint64_t)D - tile_base);
    const KeyT *kin = src ? keys1 : keys0;
    const uint32_t *vin = src ? vals1 : vals0;
    KeyT *kout = src ? keys0 : keys1;
    uint32_t *vout = src ? vals0 : vals1;

    uint32_t *wh = s_whist + warp * 256;
#pragma unroll
    for (int i = 0; i < 8; i++) wh[i * 32 + lane] = 0;

    // warp w owns the contiguous chunk [w*512, (w+1)*512) of the tile, read 32 consecutive pairs at a time, so
    // (warp, round, lane) order is the input order -- which is what makes the ranking stable.
    KeyT key[SORT_ITEMS];
    uint32_t val[SORT_ITEMS];
    const uint32_t chunk = warp * (32 * SORT_ITEMS);
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const uint32_t local = chunk + r * 32 + lane;
        const bool ok = local < n_valid;
        key[r] = ok ? __ldcs(kin + tile_base + local) : (KeyT)~(KeyT)0;
        val[r] = ok ? __ldcs(vin + tile_base + local) : 0u;
    }
    __syncwarp();
    SORT_TRACE_MARK(2)
    uint32_t rank[SORT_ITEMS];   // holds the peer mask between the two phases
    // phase 1: digit groups per round (independent across rounds: the match latencies overlap)
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const uint32_t local = chunk + r * 32 + lane;
        const uint32_t d = local < n_valid ? ((uint32_t)(key[r] >> shift) & mask) : 255u;
        rank[r] = __match_any_sync(0xffffffffu, d);   // peer mask for now
    }
    // phase 2: one shared-memory atomic per digit group and round, in round order (stable), broadcast to the group
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const uint32_t local = chunk + r * 32 + lane;
        const uint32_t d = local < n_valid ? ((uint32_t)(key[r] >> shift) & mask) : 255u;
        const uint32_t peers = rank[r];
        const uint32_t leader = 31u - (uint32_t)__clz(peers);
        uint32_t pre = 0;
        if (lane == leader) pre = atomicAdd(&wh[d], (uint32_t)__popc(peers));
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rank[r] = pre + __popc(peers & lanemask_lt());
    }
    __syncthreads();
    SORT_TRACE_MARK(3)

    // thread d (< 256): exclusive prefix of digit d over the warps, tile total, look-back for the global offset
    const uint32_t d = tid;
    uint32_t total = 0, excl = 0;
    if (tid < 256) {
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            const uint32_t c = s_whist[w * 256 + d];
            s_whist[w * 256 + d] = total;
            total += c;
        }
        const uint32_t n_invalid = SORT_TILE - n_valid;
        const uint32_t count = total - (d == 255u ? n_invalid : 0u);   // padding sorts to the very end of digit 255
        uint32_t *state = tile_state + ((size_t)pass * tiles_cap + tile) * 256;
        if (tile == 0) {
            st_volatile_u32(state + d, ST_INC | count);
        } else {
            st_volatile_u32(state + d, ST_AGG | count);
            // Look-back, eight predecessors per round trip: all CTAs of the first wave reach this point together, so
            // a one-at-a-time walk would be a chain of (resident CTAs) dependent L2 reads.
            int p = (int)tile - 1;
            bool resolved = false;
            while (!resolved) {
                uint32_t sv[LOOKBACK_WIDTH];
#pragma unroll
                for (int u = 0; u < LOOKBACK_WIDTH; u++)
                    sv[u] = p - u >= 0 ? ld_volatile_u32(state + d - (size_t)256 * (tile - (uint32_t)(p - u))) : ST_INC;
                int used = 0;
#pragma unroll
                for (int u = 0; u < LOOKBACK_WIDTH; u++) {
                    if (resolved || used != u) continue;       // stop at the first inclusive word or unready word
                    if (!st_ready(sv[u])) continue;             // not published yet: retry from here
                    excl += st_value(sv[u]);
                    used = u + 1;
                    if (st_inclusive(sv[u])) resolved = true;
                }
                p -= used;
            }
            st_volatile_u32(state + d, ST_INC | (excl + count));
        }
    }
    SORT_TRACE_MARK(4)
    const uint32_t start = block_excl_scan_256(total, s_misc, nullptr);
    if (tid < 256) {
        s_start[d] = start;
        s_gbase[d] = bins[pass * 256 + d] + excl - start;
    }
    __syncthreads();

    // reorder inside the tile
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const uint32_t local = chunk + r * 32 + lane;
        const uint32_t dg = local < n_valid ? ((uint32_t)(key[r] >> shift) & mask) : 255u;
        const uint32_t pos = s_start[dg] + s_whist[warp * 256 + dg] + rank[r];
        s_keys[pos] = key[r];
        s_vals[pos] = val[r];
    }
    __syncthreads();
    SORT_TRACE_MARK(5)
    // runs of equal digits go to consecutive global slots
    for (uint32_t i = tid; i < n_valid; i += SORT_THREADS) {
        const KeyT k = s_keys[i];
        const uint32_t dg = (uint32_t)(k >> shift) & mask;
        const uint32_t dst = s_gbase[dg] + i;
        kout[dst] = k;
        vout[dst] = s_vals[i];
    }
    SORT_TRACE_MARK(6)
}

// Depth-first binning: ranges, tile-digit histograms, pass plan and the forward's tile order from the rects' 2-D difference
// grid (rect_grid_kernel) instead of two passes over the D keys (sort_histogram_kernel before the sort, tile_ranges_kernel
// after it).  One CTA, the grid in shared memory:
//   1. prefix sums along x, then along y: cell (ty, tx) = number of duplicates of tile ty * gx + tx;
//   2. exclusive scan over the tiles in id order = where each tile's run starts in the sorted list -> ranges (a tile
//      without duplicates keeps (0, 0), like upstream's zero-initialised ranges);
//   3. digit counts of the tile passes -> bin offsets and the plan (a pass whose digit is constant is skipped);
//   4. counting sort of the tiles by list length (launch order of the blend forward).
__global__ void __launch_bounds__(256) tile_plan_kernel(RasterHeader *hdr, const int32_t *diff, int gx, int gy,
                                                        uint2 *ranges, uint32_t *bins, uint32_t *order_fwd, int sort_bits,
                                                        int num_passes, int first_pass) {
    extern __shared__ int32_t s_grid[];            // the difference grid [(gy+1)(gx+1)], then the counts [gy gx]
    __shared__ uint32_t s_hist[SORT_MAX_PASSES][256];
    __shared__ uint32_t s_tmp[8];
    __shared__ uint32_t s_bins[257];
    const int tid = threadIdx.x, pitch = gx + 1, num_tiles = gx * gy, cells = pitch * (gy + 1);
    for (int i0 = 0; i0 < cells; i0 += 256 * 8) {   // batched: a single CTA is latency-bound on these loads
        int32_t v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) { const int i = i0 + u * 256 + tid; v[u] = i < cells ? __ldcg(diff + i) : 0; }
#pragma unroll
        for (int u = 0; u < 8; u++) { const int i = i0 + u * 256 + tid; if (i < cells) s_grid[i] = v[u]; }
    }
    for (int p = 0; p < num_passes; p++) s_hist[p][tid] = 0;
    __syncthreads();
    for (int y = tid; y <= gy; y += 256) {
        int run = 0;
        for (int x = 0; x <= gx; x++) { run += s_grid[y * pitch + x]; s_grid[y * pitch + x] = run; }
    }
    __syncthreads();
    int32_t *s_cnt = s_grid + cells;               // [num_tiles] the counts in tile-id order (no div / mod later)
    for (int x = tid; x < gx; x += 256) {
        int run = 0;
        for (int y = 0; y < gy; y++) { run += s_grid[y * pitch + x]; s_cnt[y * gx + x] = run; }
    }
    __syncthreads();
    auto count_of = [&](int t) { return (uint32_t)s_cnt[t]; };
    // exclusive scan over the tiles: every thread owns a contiguous chunk
    const int per = (num_tiles + 255) / 256, t0 = min(num_tiles, tid * per), t1 = min(num_tiles, t0 + per);
    uint32_t local = 0;
    for (int t = t0; t < t1; t++) local += count_of(t);
    uint32_t total = 0;
    uint32_t start = block_excl_scan_256(local, s_tmp, &total);
    for (int t = t0; t < t1; t++) {
        const uint32_t c = count_of(t);
        ranges[t] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
        start += c;
        if (c) {
            for (int p = first_pass; p < num_passes; p++) {
                const int shift = 8 * p - 32;
                const uint32_t nb = min(8, sort_bits - 8 * p);
                atomicAdd(&s_hist[p][((uint32_t)t >> shift) & ((1u << nb) - 1u)], c);
            }
        }
    }
    __syncthreads();
    const uint32_t D = total;
    uint32_t src = 0;
    if (tid < first_pass) {   // the depth digits are already in order: those passes are skipped
        hdr->pass_skip[tid] = 1u;
        hdr->pass_src[tid] = 0u;
    }
    for (int p = first_pass; p < num_passes; p++) {
        const uint32_t c = s_hist[p][tid];
        const int all_in_one = __syncthreads_or(D == 0 || c == D);
        const uint32_t ex = block_excl_scan_256(c, s_tmp, nullptr);
        bins[p * 256 + tid] = ex;
        if (tid == 0) {
            hdr->pass_skip[p] = all_in_one ? 1u : 0u;
            hdr->pass_src[p] = src;
        }
        if (!all_in_one) src ^= 1u;
    }
    if (tid == 0) {
        hdr->final_sel = src;
        hdr->sort_bits = (uint32_t)sort_bits;
        hdr->num_passes = (uint32_t)num_passes;
    }
    __syncthreads();
    order_tiles_descending(num_tiles, count_of, order_fwd, s_bins);
}

__global__ void __launch_bounds__(256) tile_ranges_kernel(RasterHeader *hdr, const uint64_t *keys0,
                                                          const uint64_t *keys1, uint2 *ranges, int num_tiles,
                                                          uint32_t *order_fwd) {
    __shared__ uint32_t s_bins[257];
    __shared__ uint32_t s_last;
    const bool live = !hdr->overflow;
    const uint32_t D = live ? hdr->num_rendered : 0u;
    const uint64_t *keys = hdr->final_sel ? keys1 : keys0;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < D; i += stride) {
        const uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) {
            ranges[cur].x = 0;
        } else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (cur != prev) {
                ranges[prev].y = (uint32_t)i;
                ranges[cur].x = (uint32_t)i;
            }
        }
        if (i == D - 1) ranges[cur].y = D;
    }
    // the last CTA to finish orders the tiles by list length for the blend forward (longest first)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(&hdr->ranges_done, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const uint2 *r = ranges;
    order_tiles_descending(num_tiles, [r](int t) { const uint2 v = __ldcg(r + t); return v.y - v.x; }, order_fwd, s_bins);
}

}  // namespace

static int set_sort_attrs() {
    static std::atomic<uint64_t> devices{0};
    return once_per_device(devices, []() -> int {
        D3GA_CUDA(cudaFuncSetAttribute(sort_pass_kernel<uint64_t, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       sort_smem_bytes<uint64_t>()));
        D3GA_CUDA(cudaFuncSetAttribute(sort_pass_kernel<uint32_t, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       sort_smem_bytes<uint32_t>()));
        D3GA_CUDA(cudaFuncSetAttribute(tile_plan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * TILE_GRID_CELLS_MAX));
        return D3GA_OK;
    });
}

int launch_sort_and_ranges(int P, uint8_t *geom, const GeomLayout &GL, int gx, int gy, uint8_t *binning, const BinLayout &BL, uint64_t capacity, int sort_bits,
                           uint8_t *image, const ImgLayout &IL, int num_tiles, cudaStream_t stream) {
    RasterHeader *hdr = reinterpret_cast<RasterHeader *>(geom);
    uint64_t *keys0 = reinterpret_cast<uint64_t *>(binning + BL.keys0);
    uint64_t *keys1 = reinterpret_cast<uint64_t *>(binning + BL.keys1);
    uint32_t *vals0 = reinterpret_cast<uint32_t *>(binning + BL.vals0);
    uint32_t *vals1 = reinterpret_cast<uint32_t *>(binning + BL.vals1);
    uint32_t *ghist = reinterpret_cast<uint32_t *>(binning + BL.hist);
    uint32_t *bins = reinterpret_cast<uint32_t *>(binning + BL.bins);
    uint32_t *tile_state = reinterpret_cast<uint32_t *>(binning + BL.tile_state);
    uint2 *ranges = reinterpret_cast<uint2 *>(image + IL.ranges);
    const int num_passes = (sort_bits + 7) / 8;
    if (num_passes > SORT_MAX_PASSES) return D3GA_EINVAL;
    const uint32_t tiles_cap = (uint32_t)sort_tiles(capacity);
    // depth-first (default): the duplicate list was emitted in depth order, only the tile digits are left to sort.
    // global   : the full 64-bit LSD sort (depth digits first, constant digits skipped).
    const int mode = sort_mode();
    const int first_pass = mode == SORT_GLOBAL ? 0 : 4;
    // direct (default): no sort of the duplicates at all -- a stable counting scatter by tile (raster_direct.cu)
    DirectPlan direct;
    if (!recall_direct_plan(geom, P, gx, gy, capacity, &direct)) return D3GA_ENOTREADY;   // stage 1 has not run on this workspace
    if (direct.ok)
        return launch_direct_scatter(P, gx, gy, geom, GL, binning, BL, capacity, sort_bits, image, IL, direct, stream);

    {
        const int arc = set_sort_attrs();
        if (arc) return arc;
    }
    D3GA_CUDA(cudaMemsetAsync(ghist, 0, 4 * 256 * SORT_MAX_PASSES, stream));
    D3GA_CUDA(cudaMemsetAsync(tile_state, 0, 4ull * 256 * tiles_cap * num_passes, stream));
    D3GA_CUDA(cudaMemsetAsync(ranges, 0, 8ull * num_tiles, stream));

    // depth-first with a tile grid small enough for the rect difference grid: the plan comes from the per-tile counts, no
    // histogram pass over the keys before and no ranges pass after the sort
    const bool from_grid = depth_first_mode(mode) && tile_plan_from_grid(gx, gy);
    if (from_grid) {
        ProfScope prof(K_TILE_PLAN, stream);
        tile_plan_kernel<<<1, 256, 4 * ((gx + 1) * (gy + 1) + gx * gy), stream>>>(
            hdr, reinterpret_cast<const int32_t *>(geom + GL.tile_diff), gx, gy, ranges, bins,
            reinterpret_cast<uint32_t *>(image + IL.order_fwd), sort_bits, num_passes, first_pass);
    } else {
        const uint64_t per_cta = (uint64_t)HIST_THREADS * HIST_ITEMS;
        int hist_grid = (int)((capacity + per_cta - 1) / per_cta);
        if (hist_grid > 148 * 8) hist_grid = 148 * 8;
        if (hist_grid < 1) hist_grid = 1;
        ProfScope prof(K_SORT_HIST, stream);
        sort_histogram_kernel<<<hist_grid, HIST_THREADS, 0, stream>>>(hdr, keys0, ghist, bins, sort_bits, num_passes,
                                                                      first_pass);
    }
    D3GA_LAUNCH_CHECK();
    for (int p = first_pass; p < num_passes; p++) {
        const int shift = 8 * p;
        const int nb = sort_bits - shift < 8 ? sort_bits - shift : 8;
        {
            ProfScope prof(K_SORT_PASS, stream);
            sort_pass_kernel<uint64_t, 0><<<tiles_cap, SORT_THREADS, sort_smem_bytes<uint64_t>(), stream>>>(
                hdr, keys0, keys1, vals0, vals1, bins, tile_state, tiles_cap, p, shift, (1u << nb) - 1u, 0u);
        }
        D3GA_LAUNCH_CHECK();
    }
    if (from_grid) return D3GA_OK;
    int rg = (int)((capacity + 255) / 256);
    if (rg > 148 * 16) rg = 148 * 16;
    if (rg < 1) rg = 1;
    {
        ProfScope prof(K_RANGES, stream);
        tile_ranges_kernel<<<rg, 256, 0, stream>>>(hdr, keys0, keys1, ranges, num_tiles,
                                                   reinterpret_cast<uint32_t *>(image + IL.order_fwd));
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

// The 32-bit depth sort of the Gaussians that precedes the key emission in depth-first mode.
int launch_depth_sort(int P, uint8_t *geom, const GeomLayout &GL, cudaStream_t stream) {
    RasterHeader *hdr = reinterpret_cast<RasterHeader *>(geom + GL.header);
    uint32_t *dkeys = reinterpret_cast<uint32_t *>(geom + GL.dkeys);
    uint32_t *dvals = reinterpret_cast<uint32_t *>(geom + GL.dvals);
    uint32_t *bins = reinterpret_cast<uint32_t *>(geom + GL.dep_bins);
    uint32_t *state = reinterpret_cast<uint32_t *>(geom + GL.dep_state);
    const uint32_t tiles = (uint32_t)((P + SORT_TILE - 1) / SORT_TILE);
    {
        const int arc = set_sort_attrs();
        if (arc) return arc;
    }
    // (digit counts, bin offsets and the pass plan were left by preprocess_kernel)
    D3GA_CUDA(cudaMemsetAsync(state, 0, 4ull * 256 * 4 * (tiles + 1), stream));
    for (int p = 0; p < 4; p++) {
        {
            ProfScope prof(K_DEPTH_PASS, stream);
            sort_pass_kernel<uint32_t, 1><<<tiles, SORT_THREADS, sort_smem_bytes<uint32_t>(), stream>>>(
                hdr, dkeys, dkeys + GL.p_pad, dvals, dvals + GL.p_pad, bins, state, tiles, p, 8 * p, 255u, (uint32_t)P);
        }
        D3GA_LAUNCH_CHECK();
    }
    return D3GA_OK;
}

}  // namespace d3ga

#ifdef D3GA_SORT_TRACE
extern "C" int d3ga_debug_sort_trace(unsigned long long *out) {
    return cudaMemcpyFromSymbol(out, d3ga::g_sort_trace, sizeof(unsigned long long) * 4 * 512 * 8) == cudaSuccess ? 0 : -2;
}
#endif
