// raster_backward.cu -- row R8 of SURVEY.md section 8: per-Gaussian chain rule behind the blend backward.
//
// Replaces computeCov2DCUDA (backward) + preprocessCUDA (backward) of diff_gaussian_rasterization._C, reached from
// autograd when models/trainer.py:186-192 calls loss.backward() on an image produced at renderer.py:130-141.
// One thread per Gaussian, one kernel: reads the packed accumulator the blend backward filled (48 B), the mean,
// the 3-D covariance and the camera; writes every gradient tensor of the Python API directly in its AoS layout
// (no intermediate dL_dconic / dL_dcolor round trip through HBM).
//   conic -> cov2D -> (cov3D, view-space mean) ; mean2D -> mean3D through the perspective divide ;
//   optional: invdepth -> mean3D, SH colour -> (coefficients, mean3D), cov3D -> (scale, rotation).
// Algorithmic bytes per Gaussian: 48 + 12 + 24 + 4 read, 12 + 12 + 24 + 4 + 12 written = 152 B.
#include "common.cuh"

namespace d3ga {

namespace {

struct BwdParams {
    int P, sh_degree, sh_coeffs, W, H, antialiasing, has_invdepth;
    float scale_modifier, tanfovx, tanfovy, focal_x, focal_y;
    const float *means3D, *opacities, *scales, *rotations, *cov3D, *shs, *view, *proj, *campos, *grad_acc;
    const int32_t *radii;
    const uint8_t *clamped;
    const float4 *conic_opacity;
    float *dL_dmeans2D, *dL_dopacity, *dL_dcolors, *dL_dmeans3D, *dL_dcov3D, *dL_dsh, *dL_dscales, *dL_drotations;
};

__device__ __forceinline__ float tp(const float *M, int k, float x, float y, float z) {
    return M[k] * x + M[4 + k] * y + M[8 + k] * z + M[12 + k];
}

__global__ void __launch_bounds__(GEO_THREADS) preprocess_backward_kernel(const BwdParams p) {
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    const int tid = threadIdx.x;
    if (tid < 16) { s_view[tid] = p.view[tid]; s_proj[tid] = p.proj[tid]; }
    if (tid < 3) s_cam[tid] = p.campos[tid];
    __syncthreads();
    const int i = blockIdx.x * GEO_THREADS + tid;
    if (i >= p.P) return;
    float acc[GRAD_ACC_FLOATS];
    {
        const float4 *row = reinterpret_cast<const float4 *>(p.grad_acc + (size_t)i * GRAD_ACC_FLOATS);
        const float4 a = row[0], b = row[1], c = row[2];
        acc[0] = a.x; acc[1] = a.y; acc[2] = a.z; acc[3] = a.w;
        acc[4] = b.x; acc[5] = b.y; acc[6] = b.z; acc[7] = b.w;
        acc[8] = c.x; acc[9] = c.y; acc[10] = c.z; acc[11] = c.w;
    }
    // A Gaussian the blend backward never added to -- culled, or behind the surface that saturated the pixels: 86-93 % of
    // a 1 M-Gaussian avatar per camera -- has an all-zero accumulator and all-zero gradients: nothing else of it is read
    // (52 of its 104 B), nothing is computed, its zeros are written.
    bool live = false;
#pragma unroll
    for (int k = 0; k < GRAD_ACC_FLOATS; k++) live |= acc[k] != 0.f;
    live = live && p.radii[i] > 0;
    float mx = 0.f, my = 0.f, mz = 0.f, c6[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (live) {
        // the blend backward accumulates raw moments of p = G dL/dalpha about the 2-D centre:
        // (sum p dx, sum p dy, sum p dx^2, sum p dx dy, sum p dy^2, sum p); with dL/dG = opacity dL/dalpha they give
        // the six derivatives upstream accumulates per pixel
        const float4 co = p.conic_opacity[i];
        const float sx = acc[0], sy = acc[1];
        acc[0] = -co.w * (co.x * sx + co.y * sy) * (0.5f * (float)p.W);
        acc[1] = -co.w * (co.z * sy + co.y * sx) * (0.5f * (float)p.H);
        acc[2] *= -0.5f * co.w;
        acc[3] *= -0.5f * co.w;
        acc[4] *= -0.5f * co.w;
        mx = p.means3D[3 * (size_t)i]; my = p.means3D[3 * (size_t)i + 1]; mz = p.means3D[3 * (size_t)i + 2];
#pragma unroll
        for (int k = 0; k < 6; k++) c6[k] = p.cov3D[6 * (size_t)i + k];
    }
    float dmean[3] = {0.f, 0.f, 0.f}, dc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dop = live ? acc[5] : 0.f;

    if (live) {
        // ---- conic -> cov2D -> cov3D and view-space mean ------------------------------------------------------------
        float t0 = tp(s_view, 0, mx, my, mz), t1 = tp(s_view, 1, mx, my, mz);
        const float t2 = tp(s_view, 2, mx, my, mz);
        const float limx = 1.3f * p.tanfovx, limy = 1.3f * p.tanfovy;
        const float txtz = t0 / t2, tytz = t1 / t2;
        const float xm = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float ym = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        t0 = fminf(limx, fmaxf(-limx, txtz)) * t2;
        t1 = fminf(limy, fmaxf(-limy, tytz)) * t2;
        const float fx = p.focal_x, fy = p.focal_y;
        const float J00 = fx / t2, J02 = -(fx * t0) / (t2 * t2), J11 = fy / t2, J12 = -(fy * t1) / (t2 * t2);
        float Mx[3], My[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            Mx[j] = s_view[4 * j + 0] * J00 + s_view[4 * j + 2] * J02;
            My[j] = s_view[4 * j + 1] * J11 + s_view[4 * j + 2] * J12;
        }
        const float S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
        float Ux[3], Uy[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            Ux[j] = Mx[0] * S[0][j] + Mx[1] * S[1][j] + Mx[2] * S[2][j];
            Uy[j] = My[0] * S[0][j] + My[1] * S[1][j] + My[2] * S[2][j];
        }
        float cxx = Ux[0] * Mx[0] + Ux[1] * Mx[1] + Ux[2] * Mx[2];
        const float cxy = Ux[0] * My[0] + Ux[1] * My[1] + Ux[2] * My[2];
        float cyy = Uy[0] * My[0] + Uy[1] * My[1] + Uy[2] * My[2];
        float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
        if (p.antialiasing) {
            const float det_cov = cxx * cyy - cxy * cxy;
            const float pxx = cxx + 0.3f, pyy = cyy + 0.3f;
            const float det_plus = pxx * pyy - cxy * cxy;
            const float ratio = det_cov / det_plus;
            const float hs = sqrtf(fmaxf(0.000025f, ratio));
            const float dop_eff = dop;
            dop = dop_eff * hs;
            if (ratio > 0.000025f) {
                const float dr = dop_eff * p.opacities[i] * 0.5f / hs, inv = 1.f / det_plus;
                dL_da += dr * (cyy * inv - det_cov * pyy * inv * inv);
                dL_dc += dr * (cxx * inv - det_cov * pxx * inv * inv);
                dL_db += dr * (-2.f * cxy * inv + det_cov * 2.f * cxy * inv * inv);
            }
        }
        cxx += 0.3f;
        cyy += 0.3f;
        const float dcx = acc[2], dcy = acc[3], dcz = acc[4];
        const float denom = cxx * cyy - cxy * cxy;
        const float denom2inv = 1.f / (denom * denom + 0.0000001f);
        if (denom2inv != 0.f) {
            dL_da += denom2inv * (-cyy * cyy * dcx + 2.f * cxy * cyy * dcy + (denom - cxx * cyy) * dcz);
            dL_dc += denom2inv * (-cxx * cxx * dcz + 2.f * cxx * cxy * dcy + (denom - cxx * cyy) * dcx);
            dL_db += denom2inv * 2.f * (cxy * cyy * dcx - (denom + 2.f * cxy * cxy) * dcy + cxx * cxy * dcz);
        }
        dc[0] = Mx[0] * Mx[0] * dL_da + Mx[0] * My[0] * dL_db + My[0] * My[0] * dL_dc;
        dc[3] = Mx[1] * Mx[1] * dL_da + Mx[1] * My[1] * dL_db + My[1] * My[1] * dL_dc;
        dc[5] = Mx[2] * Mx[2] * dL_da + Mx[2] * My[2] * dL_db + My[2] * My[2] * dL_dc;
        dc[1] = 2.f * Mx[0] * Mx[1] * dL_da + (Mx[0] * My[1] + Mx[1] * My[0]) * dL_db + 2.f * My[0] * My[1] * dL_dc;
        dc[2] = 2.f * Mx[0] * Mx[2] * dL_da + (Mx[0] * My[2] + Mx[2] * My[0]) * dL_db + 2.f * My[0] * My[2] * dL_dc;
        dc[4] = 2.f * Mx[2] * Mx[1] * dL_da + (Mx[1] * My[2] + Mx[2] * My[1]) * dL_db + 2.f * My[1] * My[2] * dL_dc;
        float dJ00 = 0.f, dJ02 = 0.f, dJ11 = 0.f, dJ12 = 0.f;
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const float dMx = 2.f * Ux[j] * dL_da + Uy[j] * dL_db;
            const float dMy = 2.f * Uy[j] * dL_dc + Ux[j] * dL_db;
            dJ00 += s_view[4 * j + 0] * dMx; dJ02 += s_view[4 * j + 2] * dMx;
            dJ11 += s_view[4 * j + 1] * dMy; dJ12 += s_view[4 * j + 2] * dMy;
        }
        const float tz = 1.f / t2, tz2 = tz * tz, tz3 = tz2 * tz;
        const float dtx = xm * -fx * tz2 * dJ02;
        const float dty = ym * -fy * tz2 * dJ12;
        const float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2.f * fx * t0) * tz3 * dJ02 + (2.f * fy * t1) * tz3 * dJ12;
#pragma unroll
        for (int r = 0; r < 3; r++)
            dmean[r] = s_view[4 * r + 0] * dtx + s_view[4 * r + 1] * dty + s_view[4 * r + 2] * dtz;
        // ---- mean2D -> mean3D through the perspective divide ----------------------------------------------------------
        const float hw = tp(s_proj, 3, mx, my, mz);
        const float mw = 1.f / (hw + 0.0000001f);
        const float mul1 = tp(s_proj, 0, mx, my, mz) * mw * mw, mul2 = tp(s_proj, 1, mx, my, mz) * mw * mw;
#pragma unroll
        for (int r = 0; r < 3; r++)
            dmean[r] += (s_proj[4 * r + 0] * mw - s_proj[4 * r + 3] * mul1) * acc[0] +
                        (s_proj[4 * r + 1] * mw - s_proj[4 * r + 3] * mul2) * acc[1];
        if (p.has_invdepth) {
            const float zv = tp(s_view, 2, mx, my, mz);
            const float f = -acc[9] / (zv * zv);
#pragma unroll
            for (int r = 0; r < 3; r++) dmean[r] += s_view[4 * r + 2] * f;
        }
    }

    // ---- colour ----------------------------------------------------------------------------------------------------------
    if (p.shs) {
        const int M = p.sh_coeffs, deg = p.sh_degree;
        float *dsh = p.dL_dsh + (size_t)i * M * 3;
        if (!live) {
            for (int k = 0; k < M * 3; k++) dsh[k] = 0.f;
        } else {
            const float C0 = 0.28209479177387814f, C1 = 0.4886025119029199f;
            const float C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                 -1.0925484305920792f, 0.5462742152960396f};
            const float C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                 -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f};
            const float *sh = p.shs + (size_t)i * M * 3;
            const float vx = mx - s_cam[0], vy = my - s_cam[1], vz = mz - s_cam[2];
            const float s2 = vx * vx + vy * vy + vz * vz;
            const float ilen = rsqrtf(s2);
            const float x = vx * ilen, y = vy * ilen, z = vz * ilen;
            float ddx = 0.f, ddy = 0.f, ddz = 0.f;
            for (int k = (deg + 1) * (deg + 1) * 3; k < M * 3; k++) dsh[k] = 0.f;
#pragma unroll
            for (int c = 0; c < 3; c++) {
                const float g = p.clamped[3 * (size_t)i + c] ? 0.f : acc[6 + c];
                float dx_ = 0.f, dy_ = 0.f, dz_ = 0.f;
                dsh[c] = C0 * g;
                if (deg > 0) {
                    dsh[3 + c] = -C1 * y * g; dsh[6 + c] = C1 * z * g; dsh[9 + c] = -C1 * x * g;
                    dx_ = -C1 * sh[9 + c]; dy_ = -C1 * sh[3 + c]; dz_ = C1 * sh[6 + c];
                    if (deg > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        dsh[12 + c] = C2[0] * xy * g; dsh[15 + c] = C2[1] * yz * g;
                        dsh[18 + c] = C2[2] * (2.f * zz - xx - yy) * g;
                        dsh[21 + c] = C2[3] * xz * g; dsh[24 + c] = C2[4] * (xx - yy) * g;
                        dx_ += C2[0] * y * sh[12 + c] + C2[2] * 2.f * -x * sh[18 + c] + C2[3] * z * sh[21 + c] + C2[4] * 2.f * x * sh[24 + c];
                        dy_ += C2[0] * x * sh[12 + c] + C2[1] * z * sh[15 + c] + C2[2] * 2.f * -y * sh[18 + c] + C2[4] * 2.f * -y * sh[24 + c];
                        dz_ += C2[1] * y * sh[15 + c] + C2[2] * 4.f * z * sh[18 + c] + C2[3] * x * sh[21 + c];
                        if (deg > 2) {
                            dsh[27 + c] = C3[0] * y * (3.f * xx - yy) * g; dsh[30 + c] = C3[1] * xy * z * g;
                            dsh[33 + c] = C3[2] * y * (4.f * zz - xx - yy) * g;
                            dsh[36 + c] = C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * g;
                            dsh[39 + c] = C3[4] * x * (4.f * zz - xx - yy) * g; dsh[42 + c] = C3[5] * z * (xx - yy) * g;
                            dsh[45 + c] = C3[6] * x * (xx - 3.f * yy) * g;
                            dx_ += C3[0] * sh[27 + c] * 6.f * xy + C3[1] * sh[30 + c] * yz + C3[2] * sh[33 + c] * -2.f * xy +
                                   C3[3] * sh[36 + c] * -6.f * xz + C3[4] * sh[39 + c] * (-3.f * xx + 4.f * zz - yy) +
                                   C3[5] * sh[42 + c] * 2.f * xz + C3[6] * sh[45 + c] * 3.f * (xx - yy);
                            dy_ += C3[0] * sh[27 + c] * 3.f * (xx - yy) + C3[1] * sh[30 + c] * xz +
                                   C3[2] * sh[33 + c] * (-3.f * yy + 4.f * zz - xx) + C3[3] * sh[36 + c] * -6.f * yz +
                                   C3[4] * sh[39 + c] * -2.f * xy + C3[5] * sh[42 + c] * -2.f * yz + C3[6] * sh[45 + c] * -6.f * xy;
                            dz_ += C3[1] * sh[30 + c] * xy + C3[2] * sh[33 + c] * 8.f * yz +
                                   C3[3] * sh[36 + c] * 3.f * (2.f * zz - xx - yy) + C3[4] * sh[39 + c] * 8.f * xz +
                                   C3[5] * sh[42 + c] * (xx - yy);
                        }
                    }
                }
                ddx += dx_ * g; ddy += dy_ * g; ddz += dz_ * g;
            }
            const float inv32 = ilen * ilen * ilen;
            dmean[0] += ((s2 - vx * vx) * ddx - vy * vx * ddy - vz * vx * ddz) * inv32;
            dmean[1] += (-vx * vy * ddx + (s2 - vy * vy) * ddy - vz * vy * ddz) * inv32;
            dmean[2] += (-vx * vz * ddx - vy * vz * ddy + (s2 - vz * vz) * ddz) * inv32;
        }
    } else if (p.dL_dcolors) {
#pragma unroll
        for (int c = 0; c < 3; c++) p.dL_dcolors[3 * (size_t)i + c] = live ? acc[6 + c] : 0.f;
    }

    // ---- covariance --------------------------------------------------------------------------------------------------------
    if (p.scales) {
        float ds[3] = {0.f, 0.f, 0.f}, dq[4] = {0.f, 0.f, 0.f, 0.f};
        if (live) {
            const float mod = p.scale_modifier;
            const float r = p.rotations[4 * (size_t)i], x = p.rotations[4 * (size_t)i + 1],
                        y = p.rotations[4 * (size_t)i + 2], z = p.rotations[4 * (size_t)i + 3];
            const float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                                    {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                                    {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {mod * p.scales[3 * (size_t)i], mod * p.scales[3 * (size_t)i + 1], mod * p.scales[3 * (size_t)i + 2]};
            const float G[3][3] = {{dc[0], 0.5f * dc[1], 0.5f * dc[2]}, {0.5f * dc[1], dc[3], 0.5f * dc[4]}, {0.5f * dc[2], 0.5f * dc[4], dc[5]}};
            float A[3][3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                float dsk = 0.f;
#pragma unroll
                for (int a = 0; a < 3; a++) {
                    const float dLak = 2.f * (G[a][0] * Rm[0][k] + G[a][1] * Rm[1][k] + G[a][2] * Rm[2][k]) * s[k];
                    dsk += dLak * Rm[a][k];
                    A[a][k] = dLak * s[k];
                }
                ds[k] = dsk * mod;
            }
            dq[0] = 2.f * (-z * A[0][1] + y * A[0][2] + z * A[1][0] - x * A[1][2] - y * A[2][0] + x * A[2][1]);
            dq[1] = 2.f * (y * A[0][1] + z * A[0][2] + y * A[1][0] - 2.f * x * A[1][1] - r * A[1][2] + z * A[2][0] + r * A[2][1] - 2.f * x * A[2][2]);
            dq[2] = 2.f * (-2.f * y * A[0][0] + x * A[0][1] + r * A[0][2] + x * A[1][0] + z * A[1][2] - r * A[2][0] + z * A[2][1] - 2.f * y * A[2][2]);
            dq[3] = 2.f * (-2.f * z * A[0][0] - r * A[0][1] + x * A[0][2] + r * A[1][0] - 2.f * z * A[1][1] + y * A[1][2] + x * A[2][0] + y * A[2][1]);
        }
#pragma unroll
        for (int k = 0; k < 3; k++) p.dL_dscales[3 * (size_t)i + k] = ds[k];
#pragma unroll
        for (int k = 0; k < 4; k++) p.dL_drotations[4 * (size_t)i + k] = dq[k];
    }
    if (p.dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) p.dL_dcov3D[6 * (size_t)i + k] = dc[k];
    }
#pragma unroll
    for (int r = 0; r < 3; r++) p.dL_dmeans3D[3 * (size_t)i + r] = dmean[r];
    if (p.dL_dmeans2D) {
        p.dL_dmeans2D[3 * (size_t)i + 0] = live ? acc[0] : 0.f;
        p.dL_dmeans2D[3 * (size_t)i + 1] = live ? acc[1] : 0.f;
        p.dL_dmeans2D[3 * (size_t)i + 2] = 0.f;
    }
    p.dL_dopacity[i] = dop;
}

}  // namespace

int launch_preprocess_backward(int P, int sh_degree, int sh_coeffs, int W, int H, const float *means3D,
                               const float *opacities, const float *scales, float scale_modifier,
                               const float *rotations, const float *cov3D_precomp, const float *shs,
                               const float *viewmatrix, const float *projmatrix, const float *campos, float tanfovx,
                               float tanfovy, int antialiasing, const int32_t *radii, const uint8_t *geom,
                               const GeomLayout &GL, const float *grad_acc, int has_invdepth_grad,
                               float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors, float *dL_dmeans3D,
                               float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                               cudaStream_t stream) {
    if (P <= 0) return D3GA_OK;
    BwdParams p;
    p.P = P; p.sh_degree = sh_degree; p.sh_coeffs = sh_coeffs; p.W = W; p.H = H; p.antialiasing = antialiasing;
    p.has_invdepth = has_invdepth_grad;
    p.conic_opacity = reinterpret_cast<const float4 *>(geom + GL.conic_opacity);
    p.scale_modifier = scale_modifier; p.tanfovx = tanfovx; p.tanfovy = tanfovy;
    p.focal_y = (float)H / (2.0f * tanfovy);
    p.focal_x = (float)W / (2.0f * tanfovx);
    p.means3D = means3D; p.opacities = opacities; p.scales = scales; p.rotations = rotations;
    p.cov3D = cov3D_precomp ? cov3D_precomp : reinterpret_cast<const float *>(geom + GL.cov3D);
    p.shs = shs; p.view = viewmatrix; p.proj = projmatrix; p.campos = campos; p.grad_acc = grad_acc;
    p.radii = radii;
    p.clamped = geom + GL.clamped;
    p.dL_dmeans2D = dL_dmeans2D; p.dL_dopacity = dL_dopacity; p.dL_dcolors = dL_dcolors; p.dL_dmeans3D = dL_dmeans3D;
    p.dL_dcov3D = dL_dcov3D; p.dL_dsh = dL_dsh; p.dL_dscales = dL_dscales; p.dL_drotations = dL_drotations;
    ProfScope prof(K_PRE_BWD, stream);
    preprocess_backward_kernel<<<(P + GEO_THREADS - 1) / GEO_THREADS, GEO_THREADS, 0, stream>>>(p);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

}  // namespace d3ga
