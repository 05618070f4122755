/*
 * oracle/raster_oracle.c -- TEST INFRASTRUCTURE ONLY (never imported by the product path).
 *
 * CPU restatement, in plain C, of the tile-based Gaussian-splat rasterizer that D3GA calls through
 *   renderer.py:13-16,79-93,130-141  (GaussianRasterizationSettings / GaussianRasterizer).
 *
 * PARITY UNPINNED.  The algorithm lives in a third-party dependency that is ABSENT from
 * /root/reference:  graphdeco-inria/diff-gaussian-rasterization, branch dr_aa, pinned at commit
 * 9c5c2028f6fbee2be239bc4c9421ff894fe4fbe0 (.SUBMODULES.json:16-22, .gitmodules:9-12; the directory
 * submodules/diff-gaussian-rasterization/ is empty).  The reference ships no tests or golden vectors for this
 * path.  What follows restates that module's published algorithm:
 *   preprocess  : frustum cull (z_view <= 0.2), EWA projection, +0.3 low-pass, conic, radius, tile rect, SH->RGB
 *   binning     : inclusive scan of tiles_touched, (tile<<32 | depth-bits) key duplication, stable radix sort
 *                 on the low 32+bit bits, per-tile ranges
 *   blending    : front-to-back alpha compositing per 16x16 tile, alpha clamp .99, cut 1/255, stop T<1e-4
 *   backward    : back-to-front gradient of the blend, then conic/cov2D/projection/SH/cov3D chain rule
 * It is anchored on the reference's own call site (field names, argument meaning, matrix conventions of
 * lib/cameras.py:62-75 and utils/graphics_utils.py:41-75) and validated by this repo's known-answer tests and a
 * double-precision finite-difference check of the analytic backward (tests/test_oracle_raster.py).
 *
 * NUMERICS CONTRACT (the part that decides integers: depth bits, pixel centre, radius, tile rect).
 * Every operation below is an individually rounded IEEE-754 binary32 operation unless it is written FMA(a,b,c),
 * which is a single fused multiply-add.  The CUDA kernels implement the same sequence with __fmul_rn /
 * __fadd_rn / __fmaf_rn / IEEE division and sqrt, so keys, radii, rects and ranges agree bit for bit.
 * The order imitates what nvcc's default contraction produces for `m[0]*x + m[4]*y + m[8]*z + m[12]`
 * (checked with `nvcc -ptx`): t = m[4]*y; t = fma(m[0],x,t); t = fma(m[8],z,t); t = t + m[12].
 *
 * Build twice: -DREAL=float (the oracle proper) and -DREAL=double (finite-difference validation only).
 * Compile with -ffp-contract=off so the compiler never fuses on its own.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifndef REAL
#define REAL float
#endif
typedef REAL real;

#define R(x) ((real)(x))

static inline real FMA(real a, real b, real c) {
    return sizeof(real) == 4 ? (real)fmaf((float)a, (float)b, (float)c) : (real)fma((double)a, (double)b, (double)c);
}
static inline real SQRT(real a) { return sizeof(real) == 4 ? (real)sqrtf((float)a) : (real)sqrt((double)a); }
static inline real CEIL(real a) { return sizeof(real) == 4 ? (real)ceilf((float)a) : (real)ceil((double)a); }
static inline real EXP(real a) { return sizeof(real) == 4 ? (real)expf((float)a) : (real)exp((double)a); }
static inline real MAXR(real a, real b) { return a > b ? a : b; }
static inline real MINR(real a, real b) { return a < b ? a : b; }

#define TILE 16
#define NCH 3

static const real SH_C0 = R(0.28209479177387814);
static const real SH_C1 = R(0.4886025119029199);
static const real SH_C2[5] = {R(1.0925484305920792), R(-1.0925484305920792), R(0.31539156525252005),
                              R(-1.0925484305920792), R(0.5462742152960396)};
static const real SH_C3[7] = {R(-0.5900435899266435), R(2.890611442640554), R(-0.4570457994644658),
                              R(0.3731763325901154), R(-0.4570457994644658), R(1.445305721320277),
                              R(-0.5900435899266435)};

/* row-vector convention of lib/cameras.py:68-75: p_out[k] = sum_r p[r] * M[4r+k] */
static inline real tp(const real *M, int k, real x, real y, real z) {
    real t = M[4 + k] * y;
    t = FMA(M[k], x, t);
    t = FMA(M[8 + k], z, t);
    return t + M[12 + k];
}
/* dot of three products in the contraction order a1*b1 first, then fma(a0,b0), fma(a2,b2) */
static inline real dot3(real a0, real b0, real a1, real b1, real a2, real b2) {
    real t = a1 * b1;
    t = FMA(a0, b0, t);
    return FMA(a2, b2, t);
}

/* upstream ndc2Pix: evaluated in double, rounded to float at the end (same formula as lib/cameras.py:77-78) */
static inline real ndc2pix(real v, int S) { return (real)((((double)v + 1.0) * (double)S - 1.0) * 0.5); }

/* float -> int like the device's cvt.rzi.s32.f32: truncate toward zero, saturate, NaN -> 0 */
static inline int f2i(real v) {
    if (v != v) return 0;
    if (v >= R(2147483647.0)) return 2147483647;
    if (v <= R(-2147483648.0)) return (-2147483647 - 1);
    return (int)v;
}
static inline int clampi(int a, int hi) { return a < 0 ? 0 : (a > hi ? hi : a); }
static inline void get_rect(real px, real py, int radius, int gx, int gy, int *x0, int *y0, int *x1, int *y1) {
    real r = (real)radius;
    *x0 = clampi(f2i((px - r) / R(TILE)), gx);
    *y0 = clampi(f2i((py - r) / R(TILE)), gy);
    *x1 = clampi(f2i((px + r + R(TILE - 1)) / R(TILE)), gx);
    *y1 = clampi(f2i((py + r + R(TILE - 1)) / R(TILE)), gy);
}

/* Sigma = R S^2 R^T with the raw (un-normalised) quaternion (r,x,y,z), packed (xx,xy,xz,yy,yz,zz) */
static void cov3d_from_scale_rot(const real *s3, real mod, const real *q4, real *out6) {
    real r = q4[0], x = q4[1], y = q4[2], z = q4[3];
    real Rm[3][3] = {{R(1) - R(2) * FMA(y, y, z * z), R(2) * FMA(x, y, -(r * z)), R(2) * FMA(x, z, r * y)},
                     {R(2) * FMA(x, y, r * z), R(1) - R(2) * FMA(x, x, z * z), R(2) * FMA(y, z, -(r * x))},
                     {R(2) * FMA(x, z, -(r * y)), R(2) * FMA(y, z, r * x), R(1) - R(2) * FMA(x, x, y * y)}};
    real s[3] = {mod * s3[0], mod * s3[1], mod * s3[2]};
    real L[3][3];
    for (int i = 0; i < 3; i++) for (int k = 0; k < 3; k++) L[i][k] = Rm[i][k] * s[k];
    int o = 0;
    for (int i = 0; i < 3; i++) for (int j = i; j < 3; j++)
        out6[o++] = dot3(L[i][0], L[j][0], L[i][1], L[j][1], L[i][2], L[j][2]);
}

/* EWA 2-D covariance; also returns the 2x3 matrix M = Jac * Rwv and the clamp multipliers for the backward */
static void cov2d(const real *view, real fx, real fy, real tanx, real tany, const real *mean, const real *c6,
                  real *cxx, real *cxy, real *cyy, real Mx[3], real My[3], real t_out[3], real *xmul, real *ymul) {
    real t[3] = {tp(view, 0, mean[0], mean[1], mean[2]), tp(view, 1, mean[0], mean[1], mean[2]),
                 tp(view, 2, mean[0], mean[1], mean[2])};
    real limx = R(1.3) * tanx, limy = R(1.3) * tany;
    real txtz = t[0] / t[2], tytz = t[1] / t[2];
    *xmul = (txtz < -limx || txtz > limx) ? R(0) : R(1);
    *ymul = (tytz < -limy || tytz > limy) ? R(0) : R(1);
    t[0] = MINR(limx, MAXR(-limx, txtz)) * t[2];
    t[1] = MINR(limy, MAXR(-limy, tytz)) * t[2];
    real J00 = fx / t[2], J02 = -(fx * t[0]) / (t[2] * t[2]);
    real J11 = fy / t[2], J12 = -(fy * t[1]) / (t[2] * t[2]);
    for (int j = 0; j < 3; j++) {
        /* rows of the world->view rotation: r_k[j] = view[4j + k] */
        Mx[j] = FMA(view[4 * j + 2], J02, view[4 * j + 0] * J00);
        My[j] = FMA(view[4 * j + 2], J12, view[4 * j + 1] * J11);
    }
    real S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    real Ux[3], Uy[3];
    for (int j = 0; j < 3; j++) {
        Ux[j] = dot3(Mx[0], S[0][j], Mx[1], S[1][j], Mx[2], S[2][j]);
        Uy[j] = dot3(My[0], S[0][j], My[1], S[1][j], My[2], S[2][j]);
    }
    *cxx = dot3(Ux[0], Mx[0], Ux[1], Mx[1], Ux[2], Mx[2]);
    *cxy = dot3(Ux[0], My[0], Ux[1], My[1], Ux[2], My[2]);
    *cyy = dot3(Uy[0], My[0], Uy[1], My[1], Uy[2], My[2]);
    t_out[0] = t[0]; t_out[1] = t[1]; t_out[2] = t[2];
}

static void sh_to_rgb(int deg, int M, const real *mean, const real *campos, const real *sh, real *rgb, uint8_t *clamped) {
    real dx = mean[0] - campos[0], dy = mean[1] - campos[1], dz = mean[2] - campos[2];
    real len = SQRT(dx * dx + dy * dy + dz * dz);
    real x = dx / len, y = dy / len, z = dz / len;
    (void)M;
    for (int c = 0; c < 3; c++) {
#define SH(k) sh[(k) * 3 + c]
        real res = SH_C0 * SH(0);
        if (deg > 0) {
            res = res - SH_C1 * y * SH(1) + SH_C1 * z * SH(2) - SH_C1 * x * SH(3);
            if (deg > 1) {
                real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SH_C2[0] * xy * SH(4) + SH_C2[1] * yz * SH(5) + SH_C2[2] * (R(2) * zz - xx - yy) * SH(6) +
                      SH_C2[3] * xz * SH(7) + SH_C2[4] * (xx - yy) * SH(8);
                if (deg > 2) {
                    res = res + SH_C3[0] * y * (R(3) * xx - yy) * SH(9) + SH_C3[1] * xy * z * SH(10) +
                          SH_C3[2] * y * (R(4) * zz - xx - yy) * SH(11) +
                          SH_C3[3] * z * (R(2) * zz - R(3) * xx - R(3) * yy) * SH(12) +
                          SH_C3[4] * x * (R(4) * zz - xx - yy) * SH(13) + SH_C3[5] * z * (xx - yy) * SH(14) +
                          SH_C3[6] * x * (xx - R(3) * yy) * SH(15);
                }
            }
        }
#undef SH
        res += R(0.5);
        clamped[c] = res < R(0);
        rgb[c] = MAXR(res, R(0));
    }
}

/* ---- R1: preprocess forward ------------------------------------------------------------------------------- */
void orc_preprocess(int P, int deg, int M, const real *means, const real *scales, real mod, const real *rots,
                    const real *opac, const real *shs, const real *cov3D_precomp, const real *colors_precomp,
                    const real *view, const real *proj, const real *campos, int W, int H, real tanx, real tany,
                    int antialiasing,
                    int32_t *radii, real *xy, real *depths, real *cov3D, real *rgb, real *conic_opacity,
                    uint32_t *tiles_touched, uint8_t *clamped) {
    const real fy = (real)H / (R(2) * tany), fx = (real)W / (R(2) * tanx);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        radii[i] = 0; tiles_touched[i] = 0;
        xy[2 * i] = xy[2 * i + 1] = 0; depths[i] = 0;
        for (int k = 0; k < 4; k++) conic_opacity[4 * i + k] = 0;
        for (int k = 0; k < 3; k++) { clamped[3 * i + k] = 0; rgb[3 * i + k] = colors_precomp ? colors_precomp[3 * i + k] : 0; }
        const real *c6;
        if (cov3D_precomp) { c6 = cov3D_precomp + 6 * i; memcpy(cov3D + 6 * i, c6, 6 * sizeof(real)); }
        else { cov3d_from_scale_rot(scales + 3 * i, mod, rots + 4 * i, cov3D + 6 * i); c6 = cov3D + 6 * i; }
        const real *m = means + 3 * i;
        real zv = tp(view, 2, m[0], m[1], m[2]);
        if (zv <= R(0.2)) continue;
        real hx = tp(proj, 0, m[0], m[1], m[2]), hy = tp(proj, 1, m[0], m[1], m[2]), hw = tp(proj, 3, m[0], m[1], m[2]);
        real pw = R(1) / (hw + R(0.0000001));
        real px_ndc = hx * pw, py_ndc = hy * pw;
        real cxx, cxy, cyy, Mx[3], My[3], t[3], xm, ym;
        cov2d(view, fx, fy, tanx, tany, m, c6, &cxx, &cxy, &cyy, Mx, My, t, &xm, &ym);
        real det_cov = FMA(cxx, cyy, -(cxy * cxy));
        cxx += R(0.3); cyy += R(0.3);
        real det = FMA(cxx, cyy, -(cxy * cxy));
        real hscale = R(1);
        if (antialiasing) hscale = SQRT(MAXR(R(0.000025), det_cov / det));
        if (det == R(0)) continue;
        real det_inv = R(1) / det;
        real mid = R(0.5) * (cxx + cyy);
        real disc = SQRT(MAXR(R(0.1), FMA(mid, mid, -det)));
        real l1 = mid + disc, l2 = mid - disc;
        real rad = CEIL(R(3) * SQRT(MAXR(l1, l2)));
        real pxi = ndc2pix(px_ndc, W), pyi = ndc2pix(py_ndc, H);
        int x0, y0, x1, y1;
        get_rect(pxi, pyi, f2i(rad), gx, gy, &x0, &y0, &x1, &y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        if (!colors_precomp) sh_to_rgb(deg, M, m, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
        depths[i] = zv;
        radii[i] = f2i(rad);
        xy[2 * i] = pxi; xy[2 * i + 1] = pyi;
        conic_opacity[4 * i + 0] = cyy * det_inv;
        conic_opacity[4 * i + 1] = -cxy * det_inv;
        conic_opacity[4 * i + 2] = cxx * det_inv;
        conic_opacity[4 * i + 3] = opac[i] * hscale;
        tiles_touched[i] = (uint32_t)((y1 - y0) * (x1 - x0));
    }
}

/* ---- R2: inclusive scan ----------------------------------------------------------------------------------- */
uint32_t orc_inclusive_scan(int P, const uint32_t *in, uint32_t *out) {
    uint32_t s = 0;
    for (int i = 0; i < P; i++) { s += in[i]; out[i] = s; }
    return s;
}

/* ---- R3: key duplication (y outer, x inner; key = tile<<32 | bits(depth)) ---------------------------------- */
void orc_duplicate_keys(int P, const real *xy, const real *depths, const uint32_t *offsets, const int32_t *radii,
                        int W, int H, uint64_t *keys, uint32_t *vals) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        uint32_t off = i == 0 ? 0 : offsets[i - 1];
        int x0, y0, x1, y1;
        get_rect(xy[2 * i], xy[2 * i + 1], radii[i], gx, gy, &x0, &y0, &x1, &y1);
        float df = (float)depths[i];
        uint32_t dbits; memcpy(&dbits, &df, 4);
        for (int y = y0; y < y1; y++) for (int x = x0; x < x1; x++) {
            keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
            vals[off] = (uint32_t)i;
            off++;
        }
    }
}

/* ---- R4: stable LSD radix sort on the low `bits` bits ------------------------------------------------------- */
int orc_sort_bits(int W, int H) {
    uint32_t n = (uint32_t)(((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE));
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return 32 + (int)msb;
}

void orc_sort_pairs(uint64_t n, const uint64_t *kin, const uint32_t *vin, uint64_t *kout, uint32_t *vout, int bits) {
    uint64_t *ka = (uint64_t *)malloc((n + 1) * 8), *kb = (uint64_t *)malloc((n + 1) * 8);
    uint32_t *va = (uint32_t *)malloc((n + 1) * 4), *vb = (uint32_t *)malloc((n + 1) * 4);
    memcpy(ka, kin, n * 8); memcpy(va, vin, n * 4);
    for (int shift = 0; shift < bits; shift += 8) {
        int nb = bits - shift < 8 ? bits - shift : 8;
        uint64_t mask = ((uint64_t)1 << nb) - 1, cnt[257] = {0};
        for (uint64_t i = 0; i < n; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (uint64_t i = 0; i < n; i++) { uint64_t p = cnt[(ka[i] >> shift) & mask]++; kb[p] = ka[i]; vb[p] = va[i]; }
        uint64_t *tk = ka; ka = kb; kb = tk; uint32_t *tv = va; va = vb; vb = tv;
    }
    memcpy(kout, ka, n * 8); memcpy(vout, va, n * 4);
    free(ka); free(kb); free(va); free(vb);
}

/* ---- R5: per-tile [first,last) ranges ----------------------------------------------------------------------- */
void orc_tile_ranges(uint64_t L, const uint64_t *keys, int num_tiles, uint32_t *ranges /* [num_tiles,2] */) {
    memset(ranges, 0, (size_t)num_tiles * 8);
    for (uint64_t i = 0; i < L; i++) {
        uint32_t cur = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else { uint32_t prev = (uint32_t)(keys[i - 1] >> 32); if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; } }
        if (i == L - 1) ranges[2 * cur + 1] = (uint32_t)L;
    }
}

/* ---- R6: blend forward -------------------------------------------------------------------------------------- */
void orc_render_forward(int W, int H, const uint32_t *ranges, const uint32_t *point_list, const real *xy,
                        const real *features, const real *conic_opacity, const real *depths, const real *bg,
                        real *final_T, uint32_t *n_contrib, real *out_color, real *out_invdepth) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++) for (int tx = 0; tx < gx; tx++) {
        uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
        for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
            int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            real T = R(1), C[NCH] = {0, 0, 0}, invd = 0;
            uint32_t contributor = 0, last = 0;
            for (uint32_t k = r0; k < r1; k++) {
                contributor++;
                uint32_t g = point_list[k];
                real dx = xy[2 * g] - (real)px, dy = xy[2 * g + 1] - (real)py;
                const real *co = conic_opacity + 4 * g;
                real power = R(-0.5) * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                if (power > R(0)) continue;
                real alpha = MINR(R(0.99), co[3] * EXP(power));
                if (alpha < R(1) / R(255)) continue;
                real test_T = T * (R(1) - alpha);
                if (test_T < R(0.0001)) break;
                for (int c = 0; c < NCH; c++) C[c] += features[3 * g + c] * alpha * T;
                invd += (R(1) / depths[g]) * alpha * T;
                T = test_T;
                last = contributor;
            }
            size_t pid = (size_t)py * W + px;
            final_T[pid] = T; n_contrib[pid] = last;
            for (int c = 0; c < NCH; c++) out_color[(size_t)c * H * W + pid] = C[c] + T * bg[c];
            if (out_invdepth) out_invdepth[pid] = invd;
        }
    }
}

static inline void atomic_add(real *p, real v) {
#pragma omp atomic
    *p += v;
}

/* ---- R7: blend backward ------------------------------------------------------------------------------------- */
void orc_render_backward(int W, int H, const uint32_t *ranges, const uint32_t *point_list, const real *bg,
                         const real *xy, const real *conic_opacity, const real *colors, const real *depths,
                         const real *final_T, const uint32_t *n_contrib, const real *dL_dpix,
                         const real *dL_dinvdepth_pix /* nullable */,
                         real *dL_dmean2D /* [P,3] */, real *dL_dconic /* [P,4] */, real *dL_dopacity,
                         real *dL_dcolors /* [P,3] */, real *dL_dinvdepths /* [P] */) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const real ddelx_dx = R(0.5) * (real)W, ddely_dy = R(0.5) * (real)H;
#pragma omp parallel for schedule(dynamic, 1) collapse(2)
    for (int ty = 0; ty < gy; ty++) for (int tx = 0; tx < gx; tx++) {
        uint32_t r0 = ranges[2 * (ty * gx + tx)], r1 = ranges[2 * (ty * gx + tx) + 1];
        for (int ly = 0; ly < TILE; ly++) for (int lx = 0; lx < TILE; lx++) {
            int px = tx * TILE + lx, py = ty * TILE + ly;
            if (px >= W || py >= H) continue;
            size_t pid = (size_t)py * W + px;
            const real T_final = final_T[pid];
            real T = T_final;
            const uint32_t last_contributor = n_contrib[pid];
            real accum_rec[NCH] = {0, 0, 0}, last_color[NCH] = {0, 0, 0}, dLp[NCH];
            real accum_inv = 0, last_inv = 0, last_alpha = 0;
            for (int c = 0; c < NCH; c++) dLp[c] = dL_dpix[(size_t)c * H * W + pid];
            real dLinv = dL_dinvdepth_pix ? dL_dinvdepth_pix[pid] : R(0);
            real bg_dot = 0;
            for (int c = 0; c < NCH; c++) bg_dot += bg[c] * dLp[c];
            uint32_t contributor = r1 - r0;
            for (uint32_t k = r1; k-- > r0;) {
                contributor--;
                if (contributor >= last_contributor) continue;
                uint32_t g = point_list[k];
                real dx = xy[2 * g] - (real)px, dy = xy[2 * g + 1] - (real)py;
                const real *co = conic_opacity + 4 * g;
                real power = R(-0.5) * (co[0] * dx * dx + co[2] * dy * dy) - co[1] * dx * dy;
                if (power > R(0)) continue;
                real G = EXP(power);
                real alpha = MINR(R(0.99), co[3] * G);
                if (alpha < R(1) / R(255)) continue;
                T = T / (R(1) - alpha);
                real dchannel_dcolor = alpha * T;
                real dL_dalpha = 0;
                for (int c = 0; c < NCH; c++) {
                    real col = colors[3 * g + c];
                    accum_rec[c] = last_alpha * last_color[c] + (R(1) - last_alpha) * accum_rec[c];
                    last_color[c] = col;
                    dL_dalpha += (col - accum_rec[c]) * dLp[c];
                    atomic_add(&dL_dcolors[3 * g + c], dchannel_dcolor * dLp[c]);
                }
                if (dL_dinvdepth_pix) {
                    real invd = R(1) / depths[g];
                    accum_inv = last_alpha * last_inv + (R(1) - last_alpha) * accum_inv;
                    last_inv = invd;
                    dL_dalpha += (invd - accum_inv) * dLinv;
                    atomic_add(&dL_dinvdepths[g], dchannel_dcolor * dLinv);
                }
                dL_dalpha *= T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / (R(1) - alpha)) * bg_dot;
                real dL_dG = co[3] * dL_dalpha;
                real gdx = G * dx, gdy = G * dy;
                real dG_ddelx = -gdx * co[0] - gdy * co[1];
                real dG_ddely = -gdy * co[2] - gdx * co[1];
                atomic_add(&dL_dmean2D[3 * g + 0], dL_dG * dG_ddelx * ddelx_dx);
                atomic_add(&dL_dmean2D[3 * g + 1], dL_dG * dG_ddely * ddely_dy);
                atomic_add(&dL_dconic[4 * g + 0], R(-0.5) * gdx * dx * dL_dG);
                atomic_add(&dL_dconic[4 * g + 1], R(-0.5) * gdx * dy * dL_dG);
                atomic_add(&dL_dconic[4 * g + 3], R(-0.5) * gdy * dy * dL_dG);
                atomic_add(&dL_dopacity[g], G * dL_dalpha);
            }
        }
    }
}

/* ---- R8: preprocess backward (cov2D, projection, SH, cov3D chain rule) -------------------------------------- */
void orc_preprocess_backward(int P, int deg, int M, const real *means, const int32_t *radii, const real *shs,
                             const uint8_t *clamped, const real *scales, const real *rots, real mod,
                             const real *cov3Ds, const real *view, const real *proj, const real *campos,
                             int W, int H, real tanx, real tany, int antialiasing, const real *opacities,
                             const real *dL_dmean2D, const real *dL_dconic, const real *dL_dinvdepths /* nullable */,
                             real *dL_dopacity /* in/out (AA adds) */, const real *dL_dcolor,
                             real *dL_dmeans /* [P,3] out */, real *dL_dcov /* [P,6] out */,
                             real *dL_dsh /* [P,M,3] out, nullable */, real *dL_dscale /* nullable */,
                             real *dL_drot /* nullable */) {
    const real fy = (real)H / (R(2) * tany), fx = (real)W / (R(2) * tanx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        for (int k = 0; k < 3; k++) dL_dmeans[3 * i + k] = 0;
        for (int k = 0; k < 6; k++) dL_dcov[6 * i + k] = 0;
        if (dL_dsh) for (int k = 0; k < M * 3; k++) dL_dsh[(size_t)i * M * 3 + k] = 0;
        if (dL_dscale) for (int k = 0; k < 3; k++) dL_dscale[3 * i + k] = 0;
        if (dL_drot) for (int k = 0; k < 4; k++) dL_drot[4 * i + k] = 0;
        if (radii[i] <= 0) continue;
        const real *m = means + 3 * i, *c6 = cov3Ds + 6 * i;
        real cxx, cxy, cyy, Mx[3], My[3], t[3], xm, ym;
        cov2d(view, fx, fy, tanx, tany, m, c6, &cxx, &cxy, &cyy, Mx, My, t, &xm, &ym);
        real dcx = dL_dconic[4 * i + 0], dcy = dL_dconic[4 * i + 1], dcz = dL_dconic[4 * i + 3];
        real dL_dcxx = 0, dL_dcxy = 0, dL_dcyy = 0;
        if (antialiasing) {
            /* opacity' = opacity * sqrt(max(2.5e-5, det_cov / det_plus)) */
            real det_cov = cxx * cyy - cxy * cxy;
            real pxx = cxx + R(0.3), pyy = cyy + R(0.3);
            real det_plus = pxx * pyy - cxy * cxy;
            real ratio = det_cov / det_plus;
            real hs = SQRT(MAXR(R(0.000025), ratio));
            real dL_dop_eff = dL_dopacity[i];
            dL_dopacity[i] = dL_dop_eff * hs;
            if (ratio > R(0.000025)) {
                real dL_dratio = dL_dop_eff * opacities[i] * R(0.5) / hs;
                real inv = R(1) / det_plus;
                /* d ratio / d c = (d det_cov * det_plus - det_cov * d det_plus) / det_plus^2 */
                dL_dcxx += dL_dratio * (cyy * inv - det_cov * pyy * inv * inv);
                dL_dcyy += dL_dratio * (cxx * inv - det_cov * pxx * inv * inv);
                dL_dcxy += dL_dratio * (R(-2) * cxy * inv + det_cov * R(2) * cxy * inv * inv);
            }
        }
        cxx += R(0.3); cyy += R(0.3);
        real denom = cxx * cyy - cxy * cxy;
        real denom2inv = R(1) / (denom * denom + R(0.0000001));
        real dL_da = 0, dL_db = 0, dL_dc = 0;
        if (denom2inv != R(0)) {
            dL_da = denom2inv * (-cyy * cyy * dcx + R(2) * cxy * cyy * dcy + (denom - cxx * cyy) * dcz);
            dL_dc = denom2inv * (-cxx * cxx * dcz + R(2) * cxx * cxy * dcy + (denom - cxx * cyy) * dcx);
            dL_db = denom2inv * R(2) * (cxy * cyy * dcx - (denom + R(2) * cxy * cxy) * dcy + cxx * cxy * dcz);
        }
        dL_da += dL_dcxx; dL_db += dL_dcxy; dL_dc += dL_dcyy;
        real *dc = dL_dcov + 6 * i;
        dc[0] = Mx[0] * Mx[0] * dL_da + Mx[0] * My[0] * dL_db + My[0] * My[0] * dL_dc;
        dc[3] = Mx[1] * Mx[1] * dL_da + Mx[1] * My[1] * dL_db + My[1] * My[1] * dL_dc;
        dc[5] = Mx[2] * Mx[2] * dL_da + Mx[2] * My[2] * dL_db + My[2] * My[2] * dL_dc;
        dc[1] = R(2) * Mx[0] * Mx[1] * dL_da + (Mx[0] * My[1] + Mx[1] * My[0]) * dL_db + R(2) * My[0] * My[1] * dL_dc;
        dc[2] = R(2) * Mx[0] * Mx[2] * dL_da + (Mx[0] * My[2] + Mx[2] * My[0]) * dL_db + R(2) * My[0] * My[2] * dL_dc;
        dc[4] = R(2) * Mx[2] * Mx[1] * dL_da + (Mx[1] * My[2] + Mx[2] * My[1]) * dL_db + R(2) * My[1] * My[2] * dL_dc;
        real S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        real dMx[3], dMy[3];
        for (int j = 0; j < 3; j++) {
            real sx = Mx[0] * S[0][j] + Mx[1] * S[1][j] + Mx[2] * S[2][j];
            real sy = My[0] * S[0][j] + My[1] * S[1][j] + My[2] * S[2][j];
            dMx[j] = R(2) * sx * dL_da + sy * dL_db;
            dMy[j] = R(2) * sy * dL_dc + sx * dL_db;
        }
        real dJ00 = 0, dJ02 = 0, dJ11 = 0, dJ12 = 0;
        for (int j = 0; j < 3; j++) {
            dJ00 += view[4 * j + 0] * dMx[j]; dJ02 += view[4 * j + 2] * dMx[j];
            dJ11 += view[4 * j + 1] * dMy[j]; dJ12 += view[4 * j + 2] * dMy[j];
        }
        real tz = R(1) / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
        real dtx = xm * -fx * tz2 * dJ02;
        real dty = ym * -fy * tz2 * dJ12;
        real dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (R(2) * fx * t[0]) * tz3 * dJ02 + (R(2) * fy * t[1]) * tz3 * dJ12;
        real dmean[3];
        for (int r = 0; r < 3; r++) dmean[r] = view[4 * r + 0] * dtx + view[4 * r + 1] * dty + view[4 * r + 2] * dtz;
        /* projection: mean2D gradient back through the perspective divide */
        real hw = tp(proj, 3, m[0], m[1], m[2]);
        real mw = R(1) / (hw + R(0.0000001));
        real mul1 = tp(proj, 0, m[0], m[1], m[2]) * mw * mw, mul2 = tp(proj, 1, m[0], m[1], m[2]) * mw * mw;
        real gx2 = dL_dmean2D[3 * i], gy2 = dL_dmean2D[3 * i + 1];
        for (int r = 0; r < 3; r++)
            dmean[r] += (proj[4 * r + 0] * mw - proj[4 * r + 3] * mul1) * gx2 + (proj[4 * r + 1] * mw - proj[4 * r + 3] * mul2) * gy2;
        if (dL_dinvdepths) {
            real zv = tp(view, 2, m[0], m[1], m[2]);
            real f = -dL_dinvdepths[i] / (zv * zv);
            for (int r = 0; r < 3; r++) dmean[r] += view[4 * r + 2] * f;
        }
        if (shs && dL_dsh) {
            /* SH backward: colours to coefficients and to the view direction */
            real dir0[3] = {m[0] - campos[0], m[1] - campos[1], m[2] - campos[2]};
            real len = SQRT(dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2]);
            real x = dir0[0] / len, y = dir0[1] / len, z = dir0[2] / len;
            const real *sh = shs + (size_t)i * M * 3;
            real *dsh = dL_dsh + (size_t)i * M * 3;
            real ddir[3] = {0, 0, 0};
            for (int c = 0; c < 3; c++) {
                real g = clamped[3 * i + c] ? R(0) : dL_dcolor[3 * i + c];
#define SH(k) sh[(k) * 3 + c]
#define DSH(k) dsh[(k) * 3 + c]
                real dx_ = 0, dy_ = 0, dz_ = 0;
                DSH(0) = SH_C0 * g;
                if (deg > 0) {
                    DSH(1) = -SH_C1 * y * g; DSH(2) = SH_C1 * z * g; DSH(3) = -SH_C1 * x * g;
                    dx_ = -SH_C1 * SH(3); dy_ = -SH_C1 * SH(1); dz_ = SH_C1 * SH(2);
                    if (deg > 1) {
                        real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        DSH(4) = SH_C2[0] * xy * g; DSH(5) = SH_C2[1] * yz * g; DSH(6) = SH_C2[2] * (R(2) * zz - xx - yy) * g;
                        DSH(7) = SH_C2[3] * xz * g; DSH(8) = SH_C2[4] * (xx - yy) * g;
                        dx_ += SH_C2[0] * y * SH(4) + SH_C2[2] * R(2) * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * R(2) * x * SH(8);
                        dy_ += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * R(2) * -y * SH(6) + SH_C2[4] * R(2) * -y * SH(8);
                        dz_ += SH_C2[1] * y * SH(5) + SH_C2[2] * R(2) * R(2) * z * SH(6) + SH_C2[3] * x * SH(7);
                        if (deg > 2) {
                            DSH(9) = SH_C3[0] * y * (R(3) * xx - yy) * g; DSH(10) = SH_C3[1] * xy * z * g;
                            DSH(11) = SH_C3[2] * y * (R(4) * zz - xx - yy) * g;
                            DSH(12) = SH_C3[3] * z * (R(2) * zz - R(3) * xx - R(3) * yy) * g;
                            DSH(13) = SH_C3[4] * x * (R(4) * zz - xx - yy) * g; DSH(14) = SH_C3[5] * z * (xx - yy) * g;
                            DSH(15) = SH_C3[6] * x * (xx - R(3) * yy) * g;
                            dx_ += SH_C3[0] * SH(9) * R(3) * R(2) * xy + SH_C3[1] * SH(10) * yz + SH_C3[2] * SH(11) * -R(2) * xy +
                                   SH_C3[3] * SH(12) * -R(3) * R(2) * xz + SH_C3[4] * SH(13) * (-R(3) * xx + R(4) * zz - yy) +
                                   SH_C3[5] * SH(14) * R(2) * xz + SH_C3[6] * SH(15) * R(3) * (xx - yy);
                            dy_ += SH_C3[0] * SH(9) * R(3) * (xx - yy) + SH_C3[1] * SH(10) * xz + SH_C3[2] * SH(11) * (-R(3) * yy + R(4) * zz - xx) +
                                   SH_C3[3] * SH(12) * -R(3) * R(2) * yz + SH_C3[4] * SH(13) * -R(2) * xy + SH_C3[5] * SH(14) * -R(2) * yz +
                                   SH_C3[6] * SH(15) * -R(3) * R(2) * xy;
                            dz_ += SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * R(4) * R(2) * yz + SH_C3[3] * SH(12) * R(3) * (R(2) * zz - xx - yy) +
                                   SH_C3[4] * SH(13) * R(4) * R(2) * xz + SH_C3[5] * SH(14) * (xx - yy);
                        }
                    }
                }
#undef SH
#undef DSH
                ddir[0] += dx_ * g; ddir[1] += dy_ * g; ddir[2] += dz_ * g;
            }
            /* through dir = v/|v| */
            real s2 = dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2];
            real inv32 = R(1) / SQRT(s2 * s2 * s2);
            real vx = dir0[0], vy = dir0[1], vz = dir0[2];
            dmean[0] += ((s2 - vx * vx) * ddir[0] - vy * vx * ddir[1] - vz * vx * ddir[2]) * inv32;
            dmean[1] += (-vx * vy * ddir[0] + (s2 - vy * vy) * ddir[1] - vz * vy * ddir[2]) * inv32;
            dmean[2] += (-vx * vz * ddir[0] - vy * vz * ddir[1] + (s2 - vz * vz) * ddir[2]) * inv32;
        }
        for (int r = 0; r < 3; r++) dL_dmeans[3 * i + r] = dmean[r];
        if (scales && dL_dscale && dL_drot) {
            /* Sigma = L L^T, L = Rm diag(s): dL/dL = 2 G L with G the symmetric gradient (off-diagonals halved) */
            const real *q4 = rots + 4 * i;
            real r = q4[0], x = q4[1], y = q4[2], z = q4[3];
            real Rm[3][3] = {{R(1) - R(2) * (y * y + z * z), R(2) * (x * y - r * z), R(2) * (x * z + r * y)},
                             {R(2) * (x * y + r * z), R(1) - R(2) * (x * x + z * z), R(2) * (y * z - r * x)},
                             {R(2) * (x * z - r * y), R(2) * (y * z + r * x), R(1) - R(2) * (x * x + y * y)}};
            real s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
            real G[3][3] = {{dc[0], R(0.5) * dc[1], R(0.5) * dc[2]}, {R(0.5) * dc[1], dc[3], R(0.5) * dc[4]}, {R(0.5) * dc[2], R(0.5) * dc[4], dc[5]}};
            real A[3][3]; /* dL/dRm */
            for (int k = 0; k < 3; k++) {
                real ds = 0;
                for (int a = 0; a < 3; a++) {
                    real dLak = 0;
                    for (int b = 0; b < 3; b++) dLak += R(2) * G[a][b] * Rm[b][k] * s[k];
                    ds += dLak * Rm[a][k];
                    A[a][k] = dLak * s[k];
                }
                dL_dscale[3 * i + k] = ds * mod;
            }
            dL_drot[4 * i + 0] = R(2) * (-z * A[0][1] + y * A[0][2] + z * A[1][0] - x * A[1][2] - y * A[2][0] + x * A[2][1]);
            dL_drot[4 * i + 1] = R(2) * (y * A[0][1] + z * A[0][2] + y * A[1][0] - R(2) * x * A[1][1] - r * A[1][2] + z * A[2][0] + r * A[2][1] - R(2) * x * A[2][2]);
            dL_drot[4 * i + 2] = R(2) * (-R(2) * y * A[0][0] + x * A[0][1] + r * A[0][2] + x * A[1][0] + z * A[1][2] - r * A[2][0] + z * A[2][1] - R(2) * y * A[2][2]);
            dL_drot[4 * i + 3] = R(2) * (-R(2) * z * A[0][0] - r * A[0][1] + x * A[0][2] + r * A[1][0] - R(2) * z * A[1][1] + y * A[1][2] + x * A[2][0] + y * A[2][1]);
        }
    }
}

int orc_sizeof_real(void) { return (int)sizeof(real); }
