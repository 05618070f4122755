// raster_geometry.cu -- stage 1 of the rasterizer forward: rows R1 + R2 + R3 of SURVEY.md section 8 in ONE kernel.
//
//   R1 preprocess : frustum cull, EWA projection (+0.3 low-pass), conic, radius, tile rect, SH -> RGB
//   R2 scan       : prefix sum of tiles touched -- done in-kernel with a decoupled look-back over CTAs
//   R3 duplicate  : (tile << 32 | depth bits, Gaussian index) pairs, written in Gaussian order
//
// Replaces the preprocess / InclusiveSum / duplicateWithKeys sequence that
// diff_gaussian_rasterization._C.rasterize_gaussians runs when called from renderer.py:130-141.
//
// Why one kernel: R2 and R3 only need each Gaussian's tile count and rect, which R1 has in registers.  Fusing
// removes two launches, the tiles_touched/offset round trip through HBM and -- because the duplicate total lands in
// device memory and an async copy -- the mid-pipeline host synchronisation upstream needs to size its sort.
// Duplicates are emitted by the whole CTA over the CTA's flattened output range (binary search in shared memory
// for the owning Gaussian), so stores are fully coalesced and one large splat cannot serialise a warp.
//
// NUMERICS CONTRACT: everything that decides an integer (depth bits, pixel centre, radius, rect) is written with
// explicitly rounded intrinsics in the operation order documented in oracle/raster_oracle.c, so the key list is
// bit-identical to the oracle's.  Algorithmic bytes: reads 40 B (means, cov6, opacity) + 12 B (colour is read by the
// blend, not here), writes 36 B + 8 B (tiles, offsets) per Gaussian, 12 B per duplicate.
#include "common.cuh"

namespace d3ga {

namespace {

__device__ __forceinline__ float MUL(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ADD(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float SUB(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float FMA(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float DIV(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float SQRT(float a) { return __fsqrt_rn(a); }

// row-vector transform (lib/cameras.py:68-75): out[k] = sum_r p[r] * M[4r + k]
__device__ __forceinline__ float tp(const float *M, int k, float x, float y, float z) {
    float t = MUL(M[4 + k], y);
    t = FMA(M[k], x, t);
    t = FMA(M[8 + k], z, t);
    return ADD(t, M[12 + k]);
}
__device__ __forceinline__ float dot3(float a0, float b0, float a1, float b1, float a2, float b2) {
    float t = MUL(a1, b1);
    t = FMA(a0, b0, t);
    return FMA(a2, b2, t);
}
__device__ __forceinline__ float ndc2pix(float v, int S) {
    double d = __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn((double)v, 1.0), (double)S), -1.0), 0.5);
    return __double2float_rn(d);
}
__device__ __forceinline__ int clampi(int a, int hi) { return min(hi, max(0, a)); }

struct GeoParams {
    int P, sh_degree, sh_coeffs, W, H, gx, gy, antialiasing;
    float scale_modifier, tanfovx, tanfovy, focal_x, focal_y;
    const float *means3D, *scales, *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp;
    const float *view, *proj, *campos;
    RasterHeader *hdr;
    float *depth;
    float2 *xy;
    float4 *conic_opacity;
    float *rgb, *cov3D;
    uint32_t *tiles, *offsets;
    uint8_t *clamped;
    uint64_t *scan_state;
    uint64_t *keys;
    uint32_t *vals;
    uint64_t capacity;
    int32_t *radii;
    uint32_t num_ctas;
    // depth-first binning
    uint2 *rect_pack;      // [P] (x0 | y0 << 16, width) of the tile rect
    uint32_t *dkeys, *dvals;   // [2][P_pad] ping-pong of the 32-bit depth sort
    uint32_t P_pad;
    int32_t *tile_diff;    // [(gy+1)][(gx+1)] +1/-1 at the rect corners (null: grid too large, counts come from the keys)
    uint32_t *dep_ghist, *dep_bins;   // [4][256] digit counts of the depth keys / their exclusive scans (plan of the depth sort)
    uint4 *grec;           // direct tile scatter: [P] x0 | y0 << 16, width | height << 16, Gaussian id, depth bits (else null)
    SplatRecord *splat;    // [P] what the blend kernels gather per list entry (common.cuh)
    float4 *feat;          // [P] the colours (given, or from SH) as (r, g, b, 0)
};

constexpr uint64_t FLAG_AGG = 1ull << 62, FLAG_INC = 2ull << 62, FLAG_MASK = 3ull << 62;

__device__ void sh_to_rgb(int deg, const float *sh, float3 dir, float *rgb, uint8_t *clamped) {
    const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
    const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f,
                         0.5462742152960396f};
    const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                         -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
    float x = dir.x, y = dir.y, z = dir.z;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        float res = C0 * sh[c];
        if (deg > 0) {
            res = res - C1 * y * sh[3 + c] + C1 * z * sh[6 + c] - C1 * x * sh[9 + c];
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + C2[0] * xy * sh[12 + c] + C2[1] * yz * sh[15 + c] + C2[2] * (2.f * zz - xx - yy) * sh[18 + c] +
                      C2[3] * xz * sh[21 + c] + C2[4] * (xx - yy) * sh[24 + c];
                if (deg > 2) {
                    res = res + C3[0] * y * (3.f * xx - yy) * sh[27 + c] + C3[1] * xy * z * sh[30 + c] +
                          C3[2] * y * (4.f * zz - xx - yy) * sh[33 + c] +
                          C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * sh[36 + c] +
                          C3[4] * x * (4.f * zz - xx - yy) * sh[39 + c] + C3[5] * z * (xx - yy) * sh[42 + c] +
                          C3[6] * x * (xx - 3.f * yy) * sh[45 + c];
                }
            }
        }
        res += 0.5f;
        clamped[c] = res < 0.f;
        rgb[c] = fmaxf(res, 0.f);
    }
}

// ---- R1: preprocess one Gaussian; returns its tile count, rect (x0 | y0 << 16), rect width and depth bits ------------
__device__ __forceinline__ void preprocess_one(const GeoParams &p, int idx, const float *s_view, const float *s_proj,
                                               const float *s_cam, uint32_t &count, uint32_t &rect, uint32_t &width,
                                               uint32_t &dbits) {
    count = 0; rect = 0; width = 0; dbits = 0;
    {
        int radius = 0;
        float depth = 0.f, pxi = 0.f, pyi = 0.f;
        float4 co = make_float4(0.f, 0.f, 0.f, 0.f);
        const float mx = p.means3D[3 * idx], my = p.means3D[3 * idx + 1], mz = p.means3D[3 * idx + 2];
        // opacity and colour are only needed by Gaussians that survive the culls, but loading them there put a second
        // DRAM round trip behind the first (ncu: 8.4 long-scoreboard stalls per issue); they are read up front
        const float opacity_in = p.opacities[idx];
        float col_in[3] = {0.f, 0.f, 0.f};
        if (!p.shs) {
            const float *cp = p.colors_precomp + 3 * (size_t)idx;
            col_in[0] = cp[0]; col_in[1] = cp[1]; col_in[2] = cp[2];
        }
        float c6[6];
        if (p.cov3D_precomp) {
            const float2 *src = reinterpret_cast<const float2 *>(p.cov3D_precomp + 6 * (size_t)idx);
            float2 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
            c6[0] = a.x; c6[1] = a.y; c6[2] = b.x; c6[3] = b.y; c6[4] = c.x; c6[5] = c.y;
        } else {
            const float mod = p.scale_modifier;
            const float r = p.rotations[4 * idx], x = p.rotations[4 * idx + 1], y = p.rotations[4 * idx + 2],
                        z = p.rotations[4 * idx + 3];
            float Rm[3][3] = {
                {SUB(1.f, MUL(2.f, FMA(y, y, MUL(z, z)))), MUL(2.f, FMA(x, y, -MUL(r, z))), MUL(2.f, FMA(x, z, MUL(r, y)))},
                {MUL(2.f, FMA(x, y, MUL(r, z))), SUB(1.f, MUL(2.f, FMA(x, x, MUL(z, z)))), MUL(2.f, FMA(y, z, -MUL(r, x)))},
                {MUL(2.f, FMA(x, z, -MUL(r, y))), MUL(2.f, FMA(y, z, MUL(r, x))), SUB(1.f, MUL(2.f, FMA(x, x, MUL(y, y))))}};
            float s[3] = {MUL(mod, p.scales[3 * idx]), MUL(mod, p.scales[3 * idx + 1]), MUL(mod, p.scales[3 * idx + 2])};
            float L[3][3];
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int k = 0; k < 3; k++) L[i][k] = MUL(Rm[i][k], s[k]);
            int o = 0;
#pragma unroll
            for (int i = 0; i < 3; i++)
#pragma unroll
                for (int j = i; j < 3; j++) c6[o++] = dot3(L[i][0], L[j][0], L[i][1], L[j][1], L[i][2], L[j][2]);
#pragma unroll
            for (int k = 0; k < 6; k++) p.cov3D[6 * (size_t)idx + k] = c6[k];
        }
        const float zv = tp(s_view, 2, mx, my, mz);
        bool live = zv > 0.2f;
        if (live) {
            const float hx = tp(s_proj, 0, mx, my, mz), hy = tp(s_proj, 1, mx, my, mz), hw = tp(s_proj, 3, mx, my, mz);
            const float pw = DIV(1.f, ADD(hw, 0.0000001f));
            const float ndcx = MUL(hx, pw), ndcy = MUL(hy, pw);
            // EWA 2-D covariance: M = Jac * Rwv (2x3), cov = M Sigma M^T
            float t0 = tp(s_view, 0, mx, my, mz), t1 = tp(s_view, 1, mx, my, mz);
            const float t2 = zv;
            const float limx = MUL(1.3f, p.tanfovx), limy = MUL(1.3f, p.tanfovy);
            const float txtz = DIV(t0, t2), tytz = DIV(t1, t2);
            t0 = MUL(fminf(limx, fmaxf(-limx, txtz)), t2);
            t1 = MUL(fminf(limy, fmaxf(-limy, tytz)), t2);
            const float J00 = DIV(p.focal_x, t2), J02 = DIV(-MUL(p.focal_x, t0), MUL(t2, t2));
            const float J11 = DIV(p.focal_y, t2), J12 = DIV(-MUL(p.focal_y, t1), MUL(t2, t2));
            float Mx[3], My[3];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                Mx[j] = FMA(s_view[4 * j + 2], J02, MUL(s_view[4 * j + 0], J00));
                My[j] = FMA(s_view[4 * j + 2], J12, MUL(s_view[4 * j + 1], J11));
            }
            const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
            float Ux[3], Uy[3];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                Ux[j] = dot3(Mx[0], S[0][j], Mx[1], S[1][j], Mx[2], S[2][j]);
                Uy[j] = dot3(My[0], S[0][j], My[1], S[1][j], My[2], S[2][j]);
            }
            float cxx = dot3(Ux[0], Mx[0], Ux[1], Mx[1], Ux[2], Mx[2]);
            const float cxy = dot3(Ux[0], My[0], Ux[1], My[1], Ux[2], My[2]);
            float cyy = dot3(Uy[0], My[0], Uy[1], My[1], Uy[2], My[2]);
            const float det_cov = FMA(cxx, cyy, -MUL(cxy, cxy));
            cxx = ADD(cxx, 0.3f);
            cyy = ADD(cyy, 0.3f);
            const float det = FMA(cxx, cyy, -MUL(cxy, cxy));
            float hscale = 1.f;
            if (p.antialiasing) hscale = SQRT(fmaxf(0.000025f, DIV(det_cov, det)));
            live = det != 0.f;
            if (live) {
                const float det_inv = DIV(1.f, det);
                const float mid = MUL(0.5f, ADD(cxx, cyy));
                const float disc = SQRT(fmaxf(0.1f, FMA(mid, mid, -det)));
                const float l1 = ADD(mid, disc), l2 = SUB(mid, disc);
                const float rad = ceilf(MUL(3.f, SQRT(fmaxf(l1, l2))));
                pxi = ndc2pix(ndcx, p.W);
                pyi = ndc2pix(ndcy, p.H);
                const int ri = __float2int_rz(rad);
                const float rf = (float)ri;
                const int x0 = clampi(__float2int_rz(DIV(SUB(pxi, rf), 16.f)), p.gx);
                const int y0 = clampi(__float2int_rz(DIV(SUB(pyi, rf), 16.f)), p.gy);
                const int x1 = clampi(__float2int_rz(DIV(ADD(ADD(pxi, rf), 15.f), 16.f)), p.gx);
                const int y1 = clampi(__float2int_rz(DIV(ADD(ADD(pyi, rf), 15.f), 16.f)), p.gy);
                const int n = (x1 - x0) * (y1 - y0);
                live = n != 0;
                if (live) {
                    count = (uint32_t)n;
                    rect = (uint32_t)x0 | ((uint32_t)y0 << 16);
                    width = (uint32_t)(x1 - x0);
                    radius = ri;
                    depth = zv;
                    dbits = __float_as_uint(zv);
                    co = make_float4(MUL(cyy, det_inv), MUL(-cxy, det_inv), MUL(cxx, det_inv),
                                     MUL(opacity_in, hscale));
                }
            }
        }
        if (!live) { pxi = 0.f; pyi = 0.f; }
        if (live) {
            // The blend kernels' record.  The conic goes to the log2 domain (one ex2 per alpha, no extra multiply) and the
            // footprint bound with it: alpha >= 1/255 needs d^T C d <= 2 ln(255 o) =: 2 tau, i.e. q2 >= -0.5 log2(e) 2 tau,
            // taken with a relative and an absolute margin (the per-pixel test decides; this only culls).
            constexpr float LOG2E = 1.4426950408889634f;
            float lim2 = -INFINITY;                                   // never culled (degenerate conic)
            if (co.w < 1.f / 255.f) {
                lim2 = INFINITY;                                      // can never pass the cut
            } else {
                const float det = co.x * co.z - co.y * co.y, tau = __logf(255.f * co.w);
                if (det > 0.f && tau >= 0.f && co.x > 0.f && co.z > 0.f) lim2 = (-0.5f * LOG2E) * (2.f * tau * 1.002f + 0.01f);
            }
            SplatRecord r;
            r.g0 = make_float4(pxi, pyi, lim2, co.w);
            r.g1 = make_float4((-0.5f * LOG2E) * co.x, (-LOG2E) * co.y, (-0.5f * LOG2E) * co.z, DIV(1.f, depth));
            p.splat[idx] = r;
        }
        if (p.shs) {
            float rgb[3] = {0.f, 0.f, 0.f};
            uint8_t cl[3] = {0, 0, 0};
            if (live) {
                float3 d = make_float3(mx - s_cam[0], my - s_cam[1], mz - s_cam[2]);
                const float len = sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
                d.x /= len; d.y /= len; d.z /= len;
                sh_to_rgb(p.sh_degree, p.shs + (size_t)idx * p.sh_coeffs * 3, d, rgb, cl);
                p.feat[idx] = make_float4(rgb[0], rgb[1], rgb[2], 0.f);
            }
#pragma unroll
            for (int c = 0; c < 3; c++) { p.rgb[3 * (size_t)idx + c] = rgb[c]; p.clamped[3 * (size_t)idx + c] = cl[c]; }
        } else if (live) {
            p.feat[idx] = make_float4(col_in[0], col_in[1], col_in[2], 0.f);
        }
        p.depth[idx] = depth;
        p.xy[idx] = make_float2(pxi, pyi);
        p.conic_opacity[idx] = co;
        p.radii[idx] = radius;
        p.tiles[idx] = count;
    }

}

// ---- R2 + R3: CTA scan, decoupled look-back over CTAs, cooperative key emission ------------------------------------
// `slot` is this thread's position in the scan order (bid * GEO_THREADS + tid), `gidx` the Gaussian it stands for.
__device__ __forceinline__ void scan_and_emit(const GeoParams &p, uint32_t bid, int tid, bool in_range, uint32_t gidx,
                                              uint32_t count, uint32_t rect, uint32_t width, uint32_t dbits,
                                              uint32_t *s_excl, uint32_t *s_rect, uint32_t *s_width, uint32_t *s_depth,
                                              uint32_t *s_gidx, uint32_t *s_warp, uint64_t *s_base_p) {
    // ---- R2: CTA-local exclusive scan, then decoupled look-back across CTAs ---------------------------------------
    const uint32_t lane = lane_id(), warp = tid >> 5;
    uint32_t incl = count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= (uint32_t)d) incl += n;
    }
    if (lane == 31) s_warp[warp] = incl;
    s_rect[tid] = rect;
    s_width[tid] = width;
    s_depth[tid] = dbits;
    s_gidx[tid] = gidx;
    __syncthreads();
    uint32_t warp_off = 0, cta_total = 0;
#pragma unroll
    for (int w = 0; w < GEO_THREADS / 32; w++) {
        uint32_t v = s_warp[w];
        if (w < (int)warp) warp_off += v;
        cta_total += v;
    }
    const uint32_t excl = warp_off + incl - count;
    s_excl[tid] = excl;

    if (warp == 0) {
        uint64_t base = 0;
        if (bid > 0) {
            if (lane == 0) st_volatile_u64(p.scan_state + bid, FLAG_AGG | (uint64_t)cta_total);
            int look = (int)bid - 1;
            while (true) {
                const int j = look - (int)lane;
                uint64_t s = FLAG_INC;   // virtual predecessor of CTA 0: inclusive prefix 0
                if (j >= 0) {
                    do { s = ld_volatile_u64(p.scan_state + j); } while ((s & FLAG_MASK) == 0);
                }
                const uint32_t inc_mask = __ballot_sync(0xffffffffu, (s & FLAG_MASK) == FLAG_INC);
                uint64_t v = s & ~FLAG_MASK;
                if (inc_mask) {
                    const int first = __ffs(inc_mask) - 1;   // nearest predecessor holding an inclusive prefix
                    if ((int)lane > first) v = 0;
                }
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
                base += v;
                if (inc_mask) break;
                look -= 32;
            }
        }
        if (lane == 0) {
            st_volatile_u64(p.scan_state + bid, FLAG_INC | (base + cta_total));
            *s_base_p = base;
            if (bid == p.num_ctas - 1) {
                const uint64_t total = base + cta_total;
                p.hdr->num_rendered = total > 0xffffffffull ? 0xffffffffu : (uint32_t)total;
                p.hdr->overflow = total > p.capacity ? 1u : 0u;
            }
        }
    }
    __syncthreads();
    const uint64_t base = *s_base_p;
    if (in_range) p.offsets[bid * GEO_THREADS + tid] = (uint32_t)(base + excl + count);   // inclusive prefix in scan order

    // ---- R3: emit (key, value) pairs over the CTA's flattened output range ------------------------------------------
    for (uint32_t o = tid; o < cta_total; o += GEO_THREADS) {
        uint32_t g = 0;
#pragma unroll
        for (int step = GEO_THREADS / 2; step >= 1; step >>= 1)
            if (s_excl[g + step] <= o) g += step;
        const uint32_t r = o - s_excl[g], w = s_width[g], rc = s_rect[g];
        const uint32_t ty = (rc >> 16) + r / w, tx = (rc & 0xffffu) + r % w;
        const uint64_t dst = base + o;
        if (dst < p.capacity) {
            p.keys[dst] = ((uint64_t)(ty * (uint32_t)p.gx + tx) << 32) | s_depth[g];
            p.vals[dst] = s_gidx[g];
        }
    }
}

#define GEO_SMEM_DECL                                                                  \
    __shared__ uint32_t s_excl[GEO_THREADS], s_rect[GEO_THREADS], s_width[GEO_THREADS]; \
    __shared__ uint32_t s_depth[GEO_THREADS], s_gidx[GEO_THREADS];                    \
    __shared__ uint32_t s_warp[GEO_THREADS / 32];                                     \
    __shared__ uint32_t s_bid;                                                        \
    __shared__ uint64_t s_base;

// Fused R1 + R2 + R3: keys are emitted in Gaussian-index order (the order upstream's duplicateWithKeys produces).
__global__ void __launch_bounds__(GEO_THREADS) geometry_kernel(const GeoParams p) {
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    GEO_SMEM_DECL
    const int tid = threadIdx.x;
    if (tid == 0) s_bid = atomicAdd(&p.hdr->geo_ticket, 1u);
    if (tid < 16) { s_view[tid] = p.view[tid]; s_proj[tid] = p.proj[tid]; }
    if (tid < 3) s_cam[tid] = p.campos[tid];
    __syncthreads();
    const uint32_t bid = s_bid;
    const int idx = (int)(bid * GEO_THREADS + tid);
    uint32_t count = 0, rect = 0, width = 0, dbits = 0;
    if (idx < p.P) preprocess_one(p, idx, s_view, s_proj, s_cam, count, rect, width, dbits);
    scan_and_emit(p, bid, tid, idx < p.P, (uint32_t)idx, count, rect, width, dbits, s_excl, s_rect, s_width, s_depth,
                  s_gidx, s_warp, &s_base);
}

// Depth-first binning, step 1: R1 only.  Also writes the 32-bit sort key of every Gaussian (its depth bits; culled
// Gaussians get 0xffffffff) and its rect, so that the emission / scatter can run later in depth order -- and counts the
// four digits of the key while it holds it: per CTA in shared memory, the non-zero bins then added to one of
// DEP_HIST_COPIES global tables.  This used to be a kernel of its own that read the 4 N bytes of keys back (14 us at 1 M
// Gaussians, all of it launch and DRAM latency).  Measured on the way (1 M Gaussians): one global table 58 us (the top
// digits' two or three hot words serialise 3907 CTAs in the L2), 8+ tables 36 us; a last-CTA ticket for the plan needs a
// fence per CTA, which holds the CTA until its ~116 B of stores per Gaussian are acknowledged (52-60 us) -- hence no fence
// and the plan in depth_plan_kernel, behind the kernel boundary; persistent CTAs that flush once each: 40-42 us (63
// registers in the loop: four CTAs per SM instead of five).
__global__ void __launch_bounds__(GEO_THREADS, 5) preprocess_kernel(const GeoParams p) {
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    __shared__ uint32_t s_hist[4][256];
    __shared__ uint32_t s_culled;
    const int tid = threadIdx.x;
    if (tid < 16) { s_view[tid] = p.view[tid]; s_proj[tid] = p.proj[tid]; }
    if (tid < 3) s_cam[tid] = p.campos[tid];
#pragma unroll
    for (int d = 0; d < 4; d++) s_hist[d][tid] = 0;
    if (tid == 0) s_culled = 0;
    __syncthreads();
    const int idx = (int)(blockIdx.x * GEO_THREADS + tid);
    uint32_t culled = 0;
    if (idx < p.P) {
        uint32_t count, rect, width, dbits;
        preprocess_one(p, idx, s_view, s_proj, s_cam, count, rect, width, dbits);
        // the rect goes where the binning path that follows reads it: the direct tile scatter's 16-byte record, or the
        // (rect, width) pair of the depth-first emission and its rect grid
        if (p.grec) p.grec[idx] = make_uint4(rect, count ? (width | ((count / width) << 16)) : 0u, (uint32_t)idx, dbits);
        else p.rect_pack[idx] = make_uint2(rect, width);
        const uint32_t key = count ? dbits : 0xffffffffu;   // (a live Gaussian's depth is > 0.2: never all ones)
        p.dkeys[idx] = key;
        p.dvals[idx] = (uint32_t)idx;
        culled = count ? 0u : 1u;
#pragma unroll
        for (int d = 0; d < 4; d++) atomicAdd(&s_hist[d][(key >> (8 * d)) & 255u], 1u);
    }
    culled = __reduce_add_sync(0xffffffffu, culled);
    if ((tid & 31) == 0 && culled) atomicAdd(&s_culled, culled);
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 4; d++) {
        const uint32_t c = s_hist[d][tid];
        if (c) atomicAdd(&p.dep_ghist[(blockIdx.x % DEP_HIST_COPIES) * 1024 + d * 256 + tid], c);
    }
    if (tid == 0 && s_culled) atomicAdd(&p.hdr->dep_culled, s_culled);
}

// One CTA: the digit counts of the depth keys -> the plan of the depth sort (common.cuh: depth_sort_plan).
__global__ void __launch_bounds__(1024) depth_plan_kernel(RasterHeader *hdr, const uint32_t *ghist, uint32_t *bins, uint32_t n) {
    depth_sort_plan(hdr, ghist, bins, n);
}

// Per-tile duplicate counts without touching the duplicates: +1 over a Gaussian's tile rect = four corner updates of a
// 2-D difference grid whose prefix sum (tile_plan_kernel, raster_sort.cu) is the count of every tile.  The avatar covers
// a third of the image, so a few thousand cells would each take thousands of global atomics (measured: +140 us when the
// emission kernel issued them directly); a small persistent grid of CTAs therefore accumulates in shared memory and
// flushes its non-zero cells once.
__global__ void __launch_bounds__(512) rect_grid_kernel(int P, const uint2 *__restrict__ rect_pack,
                                                        const uint32_t *__restrict__ tiles, int gx, int gy,
                                                        int32_t *__restrict__ diff) {
    extern __shared__ int32_t s_grid[];
    const int pitch = gx + 1, cells = pitch * (gy + 1);
    for (int i = threadIdx.x; i < cells; i += blockDim.x) s_grid[i] = 0;
    __syncthreads();
    for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < P; g += gridDim.x * blockDim.x) {
        const uint32_t count = __ldcs(tiles + g);
        if (!count) continue;
        const uint2 rw = __ldcs(rect_pack + g);
        const int x0 = (int)(rw.x & 0xffffu), y0 = (int)(rw.x >> 16), x1 = x0 + (int)rw.y, y1 = y0 + (int)(count / rw.y);
        atomicAdd(&s_grid[y0 * pitch + x0], 1);
        atomicAdd(&s_grid[y0 * pitch + x1], -1);
        atomicAdd(&s_grid[y1 * pitch + x0], -1);
        atomicAdd(&s_grid[y1 * pitch + x1], 1);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cells; i += blockDim.x) {
        const int v = s_grid[i];
        if (v) atomicAdd(diff + i, v);
    }
}

// Depth-first binning, step 3: R2 + R3 over the Gaussians in depth order (`order` = values of the depth sort).  The
// duplicate list comes out ordered by (depth, Gaussian), so the global sort only has to run on the tile digits.
__global__ void __launch_bounds__(GEO_THREADS) emit_sorted_kernel(const GeoParams p) {
    GEO_SMEM_DECL
    const int tid = threadIdx.x;
    if (tid == 0) s_bid = atomicAdd(&p.hdr->geo_ticket, 1u);
    __syncthreads();
    const uint32_t bid = s_bid;
    const int slot = (int)(bid * GEO_THREADS + tid);
    uint32_t count = 0, rect = 0, width = 0, dbits = 0, g = 0;
    if (slot < p.P) {
        const uint32_t *order = p.hdr->dep_final_sel ? p.dvals + p.P_pad : p.dvals;
        g = order[slot];
        count = p.tiles[g];
        if (count) {
            const uint2 rw = p.rect_pack[g];
            rect = rw.x; width = rw.y;
            dbits = __float_as_uint(p.depth[g]);
        }
    }
    scan_and_emit(p, bid, tid, slot < p.P, g, count, rect, width, dbits, s_excl, s_rect, s_width, s_depth, s_gidx, s_warp,
                  &s_base);
}

// Upstream's markVisible (checkFrustum): a point is "present" when it lies in front of the near plane, z_view > 0.2 -- the
// same test, on the same arithmetic, that the preprocess culls with.
__global__ void __launch_bounds__(GEO_THREADS) mark_visible_kernel(int P, const float *__restrict__ means3D,
                                                                   const float *__restrict__ view, uint8_t *present) {
    __shared__ float s_view[16];
    if (threadIdx.x < 16) s_view[threadIdx.x] = view[threadIdx.x];
    __syncthreads();
    const int idx = blockIdx.x * GEO_THREADS + threadIdx.x;
    if (idx >= P) return;
    present[idx] = tp(s_view, 2, means3D[3 * (size_t)idx], means3D[3 * (size_t)idx + 1], means3D[3 * (size_t)idx + 2]) > 0.2f;
}

}  // namespace

int launch_mark_visible(int P, const float *means3D, const float *viewmatrix, uint8_t *present, cudaStream_t stream) {
    if (P == 0) return D3GA_OK;
    mark_visible_kernel<<<(P + GEO_THREADS - 1) / GEO_THREADS, GEO_THREADS, 0, stream>>>(P, means3D, viewmatrix, present);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

int launch_geometry(int P, int sh_degree, int sh_coeffs, const float *means3D, const float *scales,
                    float scale_modifier, const float *rotations, const float *opacities, const float *shs,
                    const float *cov3D_precomp, const float *colors_precomp, const float *viewmatrix,
                    const float *projmatrix, const float *campos, int W, int H, float tanfovx, float tanfovy,
                    int antialiasing, uint8_t *geom, const GeomLayout &GL, uint8_t *binning, const BinLayout &BL,
                    uint64_t capacity, int32_t *radii, cudaStream_t stream) {
    GeoParams p;
    p.P = P; p.sh_degree = sh_degree; p.sh_coeffs = sh_coeffs; p.W = W; p.H = H;
    p.gx = (W + TILE - 1) / TILE; p.gy = (H + TILE - 1) / TILE;
    p.antialiasing = antialiasing;
    p.scale_modifier = scale_modifier; p.tanfovx = tanfovx; p.tanfovy = tanfovy;
    p.focal_y = (float)H / (2.0f * tanfovy);
    p.focal_x = (float)W / (2.0f * tanfovx);
    p.means3D = means3D; p.scales = scales; p.rotations = rotations; p.opacities = opacities; p.shs = shs;
    p.cov3D_precomp = cov3D_precomp; p.colors_precomp = colors_precomp;
    p.view = viewmatrix; p.proj = projmatrix; p.campos = campos;
    p.hdr = reinterpret_cast<RasterHeader *>(geom + GL.header);
    p.depth = reinterpret_cast<float *>(geom + GL.depth);
    p.xy = reinterpret_cast<float2 *>(geom + GL.xy);
    p.conic_opacity = reinterpret_cast<float4 *>(geom + GL.conic_opacity);
    p.rgb = reinterpret_cast<float *>(geom + GL.rgb);
    p.cov3D = reinterpret_cast<float *>(geom + GL.cov3D);
    p.tiles = reinterpret_cast<uint32_t *>(geom + GL.tiles);
    p.offsets = reinterpret_cast<uint32_t *>(geom + GL.offsets);
    p.clamped = geom + GL.clamped;
    p.scan_state = reinterpret_cast<uint64_t *>(geom + GL.scan_state);
    p.keys = reinterpret_cast<uint64_t *>(binning + BL.keys0);
    p.vals = reinterpret_cast<uint32_t *>(binning + BL.vals0);
    p.capacity = capacity;
    p.radii = radii;
    p.num_ctas = (uint32_t)((P + GEO_THREADS - 1) / GEO_THREADS);
    // header + look-back words start at zero
    D3GA_CUDA(cudaMemsetAsync(geom + GL.header, 0, sizeof(RasterHeader), stream));
    D3GA_CUDA(cudaMemsetAsync(geom + GL.scan_state, 0, 8ull * (p.num_ctas + 1), stream));
    p.rect_pack = reinterpret_cast<uint2 *>(geom + GL.rect_pack);
    p.dkeys = reinterpret_cast<uint32_t *>(geom + GL.dkeys);
    p.dvals = reinterpret_cast<uint32_t *>(geom + GL.dvals);
    p.P_pad = (uint32_t)GL.p_pad;
    p.tile_diff = nullptr;
    const DirectPlan direct = direct_plan(P, p.gx, p.gy, capacity);
    remember_direct_plan(geom, P, p.gx, p.gy, capacity, direct, shs ? (const void *)(geom + GL.rgb) : (const void *)colors_precomp);
    p.grec = direct.ok ? reinterpret_cast<uint4 *>(geom + GL.grec) : nullptr;
    p.splat = reinterpret_cast<SplatRecord *>(geom + GL.splat);
    p.feat = reinterpret_cast<float4 *>(geom + GL.feat);
    if (depth_first_mode(sort_mode()) && !direct.ok && tile_plan_from_grid(p.gx, p.gy)) {
        p.tile_diff = reinterpret_cast<int32_t *>(geom + GL.tile_diff);
        D3GA_CUDA(cudaMemsetAsync(p.tile_diff, 0, 4ull * (p.gx + 1) * (p.gy + 1), stream));
    }
    if (p.num_ctas == 0) return D3GA_OK;
    if (!depth_first_mode(sort_mode())) {
        ProfScope prof(K_GEOMETRY, stream);
        geometry_kernel<<<p.num_ctas, GEO_THREADS, 0, stream>>>(p);
        D3GA_LAUNCH_CHECK();
        return D3GA_OK;
    }
    // depth-first binning: preprocess -> sort the Gaussians by depth (N elements) -> emit the duplicates in that order
    {
        p.dep_ghist = reinterpret_cast<uint32_t *>(geom + GL.dep_hist);
        p.dep_bins = reinterpret_cast<uint32_t *>(geom + GL.dep_bins);
        D3GA_CUDA(cudaMemsetAsync(p.dep_ghist, 0, 4 * 256 * 4 * DEP_HIST_COPIES, stream));
        const uint32_t grid = p.num_ctas;
        ProfScope prof(K_PREPROCESS, stream);
        preprocess_kernel<<<grid, GEO_THREADS, 0, stream>>>(p);
    }
    D3GA_LAUNCH_CHECK();
    {
        ProfScope prof(K_DEPTH_HIST, stream);
        depth_plan_kernel<<<1, 1024, 0, stream>>>(p.hdr, p.dep_ghist, p.dep_bins, (uint32_t)P);
    }
    D3GA_LAUNCH_CHECK();
    if (p.tile_diff) {
        ProfScope prof(K_RECT_GRID, stream);
        rect_grid_kernel<<<296, 512, 4 * (p.gx + 1) * (p.gy + 1), stream>>>(P, p.rect_pack, p.tiles, p.gx, p.gy, p.tile_diff);
        D3GA_LAUNCH_CHECK();
    }
    int rc = launch_depth_sort(P, geom, GL, stream);
    if (rc) return rc;
    // direct tile scatter: count rows instead of an emission; the duplicates are written once, in stage 2
    if (direct.ok) return launch_direct_count(P, p.gx, p.gy, geom, GL, binning, BL, capacity, direct, stream);
    {
        ProfScope prof(K_EMIT, stream);
        emit_sorted_kernel<<<p.num_ctas, GEO_THREADS, 0, stream>>>(p);
    }
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
