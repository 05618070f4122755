// raster_segsort.cu -- second half of row R4 of SURVEY.md section 8: per-tile depth sort in shared memory.
//
// The (tile|depth) keys are sorted most-significant-part first: the global onesweep passes of raster_sort.cu run on
// the TILE digits only (two 8-bit passes up to 65536 tiles), which leaves every tile's Gaussians contiguous and still
// in Gaussian-index order (the passes are stable and the unsorted list is in Gaussian order).  This kernel then sorts
// each tile's segment by its 32 depth bits with stable 8-bit passes that never leave the SM: one CTA per tile, items
// in registers, ranking by warp match-any multi-split, reorder through shared memory, constant bytes skipped per tile.
// Equal depths keep Gaussian order, so the result is bit-identical to the stable 64-bit LSD sort upstream performs
// (cub::DeviceRadixSort::SortPairs on the low 32 + bit bits) -- but three of the five global passes (24 B per
// duplicate each) become shared-memory traffic.
// Segments longer than the register/shared-memory capacity of a CTA are sorted by the same CTA with four stable
// passes that ping-pong through the two global buffers (chunk by chunk, in order), so any size is handled.
// Algorithmic bytes: 12 D read + 12 D written.
#include "common.cuh"

namespace d3ga {

namespace {

constexpr int SEG_THREADS = 512;
constexpr int SEG_WARPS = SEG_THREADS / 32;

template <int ITEMS>
struct SegSmem {
    static constexpr int CAP = SEG_THREADS * ITEMS;
    static constexpr int BYTES = CAP * 8 + 4 * (SEG_WARPS * 256 + 256 + 256 + 32);
};

// exclusive scan over the first 256 threads (all threads call; contains barriers); s_tmp: 8 words
__device__ __forceinline__ uint32_t scan256(uint32_t v, uint32_t *s_tmp) {
    const uint32_t lane = lane_id(), warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    __syncthreads();
    if (lane == 31 && warp < 8) s_tmp[warp] = incl;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < 8; w++)
        if (w < (int)warp) off += s_tmp[w];
    return off + incl - v;
}

// One stable counting pass over the `n` items a CTA holds in registers (warp w owns the contiguous index range
// [w * per_warp, (w+1) * per_warp), round r covers 32 consecutive items): fills rank[] and the shared tables so that
// item (w, r, lane) belongs at  s_start[digit] + s_whist[w][digit] + rank[r].
template <int ITEMS>
__device__ __forceinline__ void rank_items(const uint32_t (&dep)[ITEMS], uint32_t (&rank)[ITEMS], int n, int per_warp,
                                           int rounds, int shift, uint32_t *s_whist, uint32_t *s_start,
                                           uint32_t *s_total, uint32_t *s_tmp) {
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    uint32_t *wh = s_whist + warp * 256;
#pragma unroll
    for (int i = 0; i < 8; i++) wh[i * 32 + lane] = 0;
    __syncwarp();
    // phase 1: who shares my digit in this round?  Independent across rounds, so the match latencies overlap.
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        if (r < rounds) {
            const int idx = (int)warp * per_warp + r * 32 + (int)lane;
            const uint32_t d = idx < n ? ((dep[r] >> shift) & 255u) : 255u;   // padding ranks behind every real item
            rank[r] = match_digit8(d);   // peer mask for now
        }
    }
    // phase 2: one shared-memory atomic per digit group and round (in round order: stable), broadcast to the group
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        if (r < rounds) {
            const int idx = (int)warp * per_warp + r * 32 + (int)lane;
            const uint32_t d = idx < n ? ((dep[r] >> shift) & 255u) : 255u;
            const uint32_t peers = rank[r];
            const uint32_t leader = 31u - (uint32_t)__clz(peers);
            uint32_t pre = 0;
            if (lane == leader) pre = atomicAdd(&wh[d], (uint32_t)__popc(peers));
            pre = __shfl_sync(0xffffffffu, pre, leader);
            rank[r] = pre + __popc(peers & lanemask_lt());
        }
    }
    __syncthreads();
    uint32_t total = 0;
    if (tid < 256) {
#pragma unroll
        for (int w = 0; w < SEG_WARPS; w++) {
            const uint32_t c = s_whist[w * 256 + tid];
            s_whist[w * 256 + tid] = total;
            total += c;
        }
        s_total[tid] = total;
    }
    const uint32_t start = scan256(total, s_tmp);
    if (tid < 256) s_start[tid] = start;
    __syncthreads();
}

template <int ITEMS>
__global__ void __launch_bounds__(SEG_THREADS) segment_sort_kernel(const RasterHeader *__restrict__ hdr,
                                                                   uint64_t *keys0, uint64_t *keys1, uint32_t *vals0,
                                                                   uint32_t *vals1, const uint2 *__restrict__ ranges,
                                                                   const uint32_t *__restrict__ order, int num_tiles,
                                                                   uint32_t min_len, uint32_t max_len) {
    extern __shared__ __align__(16) uint8_t smem[];
    constexpr int CAP = SegSmem<ITEMS>::CAP;
    uint32_t *s_dep = reinterpret_cast<uint32_t *>(smem);
    uint32_t *s_val = s_dep + CAP;
    uint32_t *s_whist = s_val + CAP;
    uint32_t *s_start = s_whist + SEG_WARPS * 256;
    uint32_t *s_total = s_start + 256;
    uint32_t *s_tmp = s_total + 256;   // [0..7] scan, [8] and-reduce, [9] or-reduce

    if (hdr->overflow) return;
    const uint32_t tile = order[blockIdx.x];
    const uint2 range = ranges[tile];
    const uint32_t len = range.y - range.x;
    if (len < min_len || len > max_len || len < 2) return;
    const int tid = threadIdx.x;
    const uint32_t lane = tid & 31, warp = tid >> 5;
    uint64_t *kmain = hdr->final_sel ? keys1 : keys0, *kalt = hdr->final_sel ? keys0 : keys1;
    uint32_t *vmain = hdr->final_sel ? vals1 : vals0, *valt = hdr->final_sel ? vals0 : vals1;
    const uint64_t tile_bits = (uint64_t)tile << 32;

    uint32_t dep[ITEMS], val[ITEMS], rank[ITEMS];
    if (len <= (uint32_t)CAP) {
        // ---- the whole segment lives in this CTA: registers <-> shared memory, HBM touched twice -----------------
        const int n = (int)len;
        const int per_warp = ((n + SEG_WARPS - 1) / SEG_WARPS + 31) & ~31;
        const int rounds = per_warp / 32;
        uint32_t a = 0xffffffffu, o = 0u;
#pragma unroll
        for (int r = 0; r < ITEMS; r++) {
            dep[r] = 0; val[r] = 0;
            if (r < rounds) {
                const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                if (idx < n) {
                    dep[r] = (uint32_t)kmain[range.x + idx];
                    val[r] = vmain[range.x + idx];
                    a &= dep[r]; o |= dep[r];
                }
            }
        }
        if (tid == 0) { s_tmp[8] = 0xffffffffu; s_tmp[9] = 0u; }
        __syncthreads();
        a = __reduce_and_sync(0xffffffffu, a);
        o = __reduce_or_sync(0xffffffffu, o);
        if (lane == 0) { atomicAnd(&s_tmp[8], a); atomicOr(&s_tmp[9], o); }
        __syncthreads();
        const uint32_t varying = s_tmp[8] ^ s_tmp[9];   // bits that differ somewhere in the segment
        for (int shift = 0; shift < 32; shift += 8) {
            if (((varying >> shift) & 255u) == 0) continue;   // constant byte (e.g. the float exponent): nothing to do
            rank_items<ITEMS>(dep, rank, n, per_warp, rounds, shift, s_whist, s_start, s_total, s_tmp);
#pragma unroll
            for (int r = 0; r < ITEMS; r++) {
                if (r < rounds) {
                    const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                    if (idx < n) {
                        const uint32_t d = (dep[r] >> shift) & 255u;
                        const uint32_t pos = s_start[d] + s_whist[warp * 256 + d] + rank[r];
                        s_dep[pos] = dep[r];
                        s_val[pos] = val[r];
                    }
                }
            }
            __syncthreads();
#pragma unroll
            for (int r = 0; r < ITEMS; r++) {
                if (r < rounds) {
                    const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                    if (idx < n) { dep[r] = s_dep[idx]; val[r] = s_val[idx]; }
                }
            }
            // the next pass's scatter is separated from these reads by the barriers inside rank_items
        }
#pragma unroll
        for (int r = 0; r < ITEMS; r++) {
            if (r < rounds) {
                const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                if (idx < n) {
                    kmain[range.x + idx] = tile_bits | dep[r];
                    vmain[range.x + idx] = val[r];
                }
            }
        }
        return;
    }
    // ---- oversized segment: four stable passes through global memory, chunk by chunk, by this one CTA -------------
    uint32_t *s_base = s_dep;   // [256] running global offset of each digit (the staging area is unused here)
    for (int pass = 0; pass < 4; pass++) {
        const int shift = 8 * pass;
        const uint64_t *ksrc = (pass & 1) ? kalt : kmain;
        const uint32_t *vsrc = (pass & 1) ? valt : vmain;
        uint64_t *kdst = (pass & 1) ? kmain : kalt;
        uint32_t *vdst = (pass & 1) ? vmain : valt;
        __syncthreads();
        if (tid < 256) s_total[tid] = 0;
        __syncthreads();
        for (uint32_t i = tid; i < len; i += SEG_THREADS)
            atomicAdd(&s_total[((uint32_t)ksrc[range.x + i] >> shift) & 255u], 1u);
        __syncthreads();
        const uint32_t cnt = tid < 256 ? s_total[tid] : 0u;
        const uint32_t ex = scan256(cnt, s_tmp);
        if (tid < 256) s_base[tid] = ex;
        __syncthreads();
        for (uint32_t c0 = 0; c0 < len; c0 += CAP) {
            const int n = (int)min((uint32_t)CAP, len - c0);
            const int per_warp = ((n + SEG_WARPS - 1) / SEG_WARPS + 31) & ~31;
            const int rounds = per_warp / 32;
#pragma unroll
            for (int r = 0; r < ITEMS; r++) {
                dep[r] = 0; val[r] = 0;
                if (r < rounds) {
                    const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                    if (idx < n) {
                        dep[r] = (uint32_t)ksrc[range.x + c0 + idx];
                        val[r] = vsrc[range.x + c0 + idx];
                    }
                }
            }
            rank_items<ITEMS>(dep, rank, n, per_warp, rounds, shift, s_whist, s_start, s_total, s_tmp);
#pragma unroll
            for (int r = 0; r < ITEMS; r++) {
                if (r < rounds) {
                    const int idx = (int)warp * per_warp + r * 32 + (int)lane;
                    if (idx < n) {
                        const uint32_t d = (dep[r] >> shift) & 255u;
                        const uint32_t pos = s_base[d] + s_whist[warp * 256 + d] + rank[r];
                        kdst[range.x + pos] = tile_bits | dep[r];
                        vdst[range.x + pos] = val[r];
                    }
                }
            }
            __syncthreads();
            if (tid < 256) s_base[tid] += s_total[tid] - (tid == 255 ? (uint32_t)(SEG_WARPS * per_warp - n) : 0u);
            __syncthreads();
        }
    }
}

}  // namespace

int launch_segment_sort(uint8_t *geom, uint8_t *binning, const BinLayout &BL, uint8_t *image, const ImgLayout &IL,
                        int num_tiles, cudaStream_t stream) {
    const RasterHeader *hdr = reinterpret_cast<const RasterHeader *>(geom);
    uint64_t *keys0 = reinterpret_cast<uint64_t *>(binning + BL.keys0);
    uint64_t *keys1 = reinterpret_cast<uint64_t *>(binning + BL.keys1);
    uint32_t *vals0 = reinterpret_cast<uint32_t *>(binning + BL.vals0);
    uint32_t *vals1 = reinterpret_cast<uint32_t *>(binning + BL.vals1);
    const uint2 *ranges = reinterpret_cast<const uint2 *>(image + IL.ranges);
    const uint32_t *order = reinterpret_cast<const uint32_t *>(image + IL.order_fwd);
    static bool attr = false;
    if (!attr) {
        D3GA_CUDA(cudaFuncSetAttribute(segment_sort_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SegSmem<8>::BYTES));
        D3GA_CUDA(cudaFuncSetAttribute(segment_sort_kernel<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, SegSmem<24>::BYTES));
        attr = true;
    }
    // two size classes (register budget): long segments first, they are also first in `order`
    {
        ProfScope prof(K_SEG_SORT, stream);
        segment_sort_kernel<24><<<num_tiles, SEG_THREADS, SegSmem<24>::BYTES, stream>>>(
            hdr, keys0, keys1, vals0, vals1, ranges, order, num_tiles, (uint32_t)SegSmem<8>::CAP + 1u, 0xffffffffu);
    }
    D3GA_LAUNCH_CHECK();
    {
        ProfScope prof(K_SEG_SORT, stream);
        segment_sort_kernel<8><<<num_tiles, SEG_THREADS, SegSmem<8>::BYTES, stream>>>(
            hdr, keys0, keys1, vals0, vals1, ranges, order, num_tiles, 2u, (uint32_t)SegSmem<8>::CAP);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
