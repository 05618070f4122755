// deform.cu -- rows A1 / A2 of SURVEY.md section 8: cage -> Gaussian deformation, forward and backward, one kernel each.
//
// Replaces the ~25-launch torch chain of models/cage_net.py:218-230:
//   deformed_tetras = tetpoints[tetra_faces]                               cage_net.py:218
//   J   = G(deformed_tetras[tetra_id]) @ canonical_gradient                lib/cage.py:339-342, tetra.py:74-80
//   cov = J (L L^T) J^T,  L = R(q / |q|) diag(s)                           cage_net.py:161-164,221-222,
//                                                                          utils/general_utils.py:58-90
//   cov3D_precomp = (xx, xy, xz, yy, yz, zz)                               utils/general_utils.py:24-34
//   means3D = sum_k barys[:, k] * deformed_tetras[tetra_id][:, k]          cage_net.py:228-230
// and its autograd backward (index_put_ with accumulate into tetpoints).
// One thread per Gaussian; nothing but inputs and outputs touches HBM (84 B read + 36 B written per Gaussian forward
// against >= 0.6 kB of temporaries in the torch chain).  The cage (<= a few hundred kB) stays L2 resident.
//
// The same launches optionally take in what models/cage_net.py computes right before and beside that chain (the fused
// entry points, d3ga_deform_forward_fused / _backward_fused):
//   canon_barys = barys + delta_bary                  cage_net.py:213   (not renormalised)
//   scales      = exp(scaling + delta_scale)          cage_net.py:214   (scaling_activation = exp, :78)
//   rotation    = F.normalize(rotation + delta_rot)   cage_net.py:215   (the kernel normalises once: build_rotation's own
//                                                                        normalisation makes F.normalize's a no-op except
//                                                                        for the zero vector, NaN in both)
//   fm_energy   = fem_energy(tetpoints)               cage_net.py:225, lib/cage.py:349-361 -- per-TET work, run by extra
//                                                      CTAs appended to the grid (it gathers the same tetpoints[tets])
// Base tensor and delta receive the same gradient, which the backward writes once.
#include "common.cuh"
#include "fem.cuh"

namespace d3ga {

namespace {

#ifndef DEFORM_BWD_MIN_CTAS
#define DEFORM_BWD_MIN_CTAS 3
#endif

struct Tet4 {
    float v[4][3];
};

// What the fused entry points add to a launch; all null / zero for the plain ones.
struct DeformExtra {
    const float *delta_bary, *delta_scale, *delta_rot;   // [N,4], [N,3], [N,4] or null
    int log_scales;                                      // scales are pre-activation: s = exp(scales (+ delta_scale))
    const float *dn_inv;                                 // [T,3,3] inverse rest-pose edge matrices, or null (no energy)
    float *energy;                                       // forward: [T] out
    const float *dL_denergy;                             // backward: [T] or null
    int gaussian_blocks;                                 // CTAs below this index work on Gaussians, the rest on tets
    int grad_pitch;                                      // floats per row of dL_dtetpoints: 3, or 4 (padded: vector REDs)
    int tet_sorted;                                      // rows are already in tet order: warp-summed like `order`
    const uint32_t *count;                               // backward: `order` holds *count entries (the sparse form), or null
    int plain_list;                                      // backward: `order` is no tet order: no warp sums
};

// barys, scales (activated) and the un-normalised quaternion of Gaussian n as the kernels use them; `sraw` = the scale
// row as it lies in memory (Delta already added), read by the caller
__device__ __forceinline__ void load_gaussian(const DeformExtra &ex, const float *__restrict__ barys,
                                              const float (&sraw)[3], const float *__restrict__ rotation,
                                              size_t n, float (&bw)[4], float (&s)[3], float4 &q) {
    float4 b = *reinterpret_cast<const float4 *>(barys + 4 * n);
    if (ex.delta_bary) {
        const float4 d = *reinterpret_cast<const float4 *>(ex.delta_bary + 4 * n);
        b.x += d.x; b.y += d.y; b.z += d.z; b.w += d.w;
    }
    bw[0] = b.x; bw[1] = b.y; bw[2] = b.z; bw[3] = b.w;
#pragma unroll
    for (int k = 0; k < 3; k++) s[k] = ex.log_scales ? expf(sraw[k]) : sraw[k];
    q = *reinterpret_cast<const float4 *>(rotation + 4 * n);
    if (ex.delta_rot) {
        const float4 d = *reinterpret_cast<const float4 *>(ex.delta_rot + 4 * n);
        q.x += d.x; q.y += d.y; q.z += d.z; q.w += d.w;
    }
}

__device__ __forceinline__ void load_tet(const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
                                         int64_t t, int64_t (&vid)[4], Tet4 &X) {
    const longlong2 *row = reinterpret_cast<const longlong2 *>(tets + 4 * t);   // one 32-byte row: two 16-byte loads
    const longlong2 a = __ldg(row), b = __ldg(row + 1);
    vid[0] = a.x; vid[1] = a.y; vid[2] = b.x; vid[3] = b.y;
#pragma unroll
    for (int k = 0; k < 4; k++)
#pragma unroll
        for (int c = 0; c < 3; c++) X.v[k][c] = __ldg(tetpoints + 3 * vid[k] + c);
}

__device__ __forceinline__ void rotation_from_unit_quat(float r, float x, float y, float z, float (&Rm)[3][3]) {
    Rm[0][0] = 1.f - 2.f * (y * y + z * z); Rm[0][1] = 2.f * (x * y - r * z); Rm[0][2] = 2.f * (x * z + r * y);
    Rm[1][0] = 2.f * (x * y + r * z); Rm[1][1] = 1.f - 2.f * (x * x + z * z); Rm[1][2] = 2.f * (y * z - r * x);
    Rm[2][0] = 2.f * (x * z - r * y); Rm[2][1] = 2.f * (y * z + r * x); Rm[2][2] = 1.f - 2.f * (x * x + y * y);
}

__global__ void __launch_bounds__(GEO_THREADS) deform_forward_kernel(
    int N, int T, const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
    const int64_t *__restrict__ tetra_id, const float *__restrict__ barys, const float *__restrict__ cgrad,
    const float *__restrict__ scales, const float *__restrict__ rotation, float *__restrict__ means3D,
    float *__restrict__ cov3D, const DeformExtra ex) {
    if ((int)blockIdx.x >= ex.gaussian_blocks) {   // appended CTAs: the cage's neo-Hookean energy, one thread per tet
        const int tt = ((int)blockIdx.x - ex.gaussian_blocks) * GEO_THREADS + (int)threadIdx.x;
        if (tt < T) fem::fem_forward_tet(tt, tetpoints, tets, ex.dn_inv, ex.energy);
        return;
    }
    __shared__ float s_stage[GEO_THREADS / 32][STAGE_FLOATS];
    const uint32_t lane = threadIdx.x & 31;
    float *s_w = s_stage[threadIdx.x >> 5];
    const size_t row0 = (size_t)blockIdx.x * GEO_THREADS + (threadIdx.x & ~31u);   // the warp's first Gaussian
    const bool live = row0 + lane < (size_t)N;
    const int n = live ? (int)(row0 + lane) : N - 1;          // dead lanes of the last warp shadow the last Gaussian
    // the [N,K] streams, coalesced through shared memory
    float cg9[9], sraw[3];
    int64_t t = __ldg(tetra_id + n);                         // head of the only dependent chain: id -> tet -> vertices
    {
        float r9[9], r3[3], rd[3] = {0.f, 0.f, 0.f};
        warp_rows_fetch<9>(cgrad, row0, (size_t)N, lane, r9);
        warp_rows_fetch<3>(scales, row0, (size_t)N, lane, r3);
        if (ex.delta_scale) warp_rows_fetch<3>(ex.delta_scale, row0, (size_t)N, lane, rd);   // same offsets: added raw
        r3[0] += rd[0]; r3[1] += rd[1]; r3[2] += rd[2];
        warp_rows_take<9>(s_w, lane, r9, cg9);
        warp_rows_take<3>(s_w, lane, r3, sraw);
    }
    if (t < 0) t += T;   // torch indexing wraps a -1 from compute_bary to the last tet
    float bw[4], s[3];
    float4 q;
    load_gaussian(ex, barys, sraw, rotation, (size_t)n, bw, s, q);
    int64_t vid[4];
    Tet4 X;
    load_tet(tetpoints, tets, t, vid, X);
    float mean[3];
#pragma unroll
    for (int c = 0; c < 3; c++) mean[c] = X.v[0][c] * bw[0] + X.v[1][c] * bw[1] + X.v[2][c] * bw[2] + X.v[3][c] * bw[3];
    warp_rows_store<3>(means3D, row0, (size_t)N, s_w, lane, mean);
    // G: edges (v3-v0, v2-v0, v1-v0) as columns
    float G[3][3], Cg[3][3], J[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        G[i][0] = X.v[3][i] - X.v[0][i]; G[i][1] = X.v[2][i] - X.v[0][i]; G[i][2] = X.v[1][i] - X.v[0][i];
    }
#pragma unroll
    for (int k = 0; k < 9; k++) Cg[k / 3][k % 3] = cg9[k];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) J[i][j] = G[i][0] * Cg[0][j] + G[i][1] * Cg[1][j] + G[i][2] * Cg[2][j];
    const float inv = 1.f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    float Rm[3][3];
    rotation_from_unit_quat(q.x * inv, q.y * inv, q.z * inv, q.w * inv, Rm);
    float B[3][3];   // B = J R diag(s); cov = B B^T
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int k = 0; k < 3; k++) B[i][k] = (J[i][0] * Rm[0][k] + J[i][1] * Rm[1][k] + J[i][2] * Rm[2][k]) * s[k];
    float c6[6];
    int o = 0;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = i; j < 3; j++) c6[o++] = B[i][0] * B[j][0] + B[i][1] * B[j][1] + B[i][2] * B[j][2];
    warp_rows_store<6>(cov3D, row0, (size_t)N, s_w, lane, c6);
}

// Warp-level segmented sum over runs of equal tet id (lanes are tet-sorted), then one set of 12 atomics per run.
// Every lane of the warp calls it from the same place (full-mask shuffles); lanes without a contribution pass zeros.
__device__ __forceinline__ void segmented_vertex_scatter(int64_t t, const int64_t (&vid)[4], float (&v)[12],
                                                         float *__restrict__ dL_dtetpoints, int pitch) {
    const uint32_t lane = threadIdx.x & 31;
    const int64_t tprev = __shfl_up_sync(0xffffffffu, t, 1);
    const bool head = lane == 0 || t != tprev;
    const uint32_t heads = __ballot_sync(0xffffffffu, head);
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        // lane may absorb lane+off iff no run starts in (lane, lane+off]
        const bool take = lane + off < 32 && ((heads >> (lane + 1)) & ((1u << off) - 1u)) == 0u;
#pragma unroll
        for (int k = 0; k < 12; k++) {
            const float o = __shfl_down_sync(0xffffffffu, v[k], off);
            if (take) v[k] += o;
        }
    }
    if (head) {
        if (pitch == 4) {   // padded rows: one 16-byte vector reduction per vertex
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (v[3 * k] != 0.f || v[3 * k + 1] != 0.f || v[3 * k + 2] != 0.f)
                    red_add_v4(dL_dtetpoints + 4 * vid[k], v[3 * k], v[3 * k + 1], v[3 * k + 2], 0.f);
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int c = 0; c < 3; c++)
                    if (v[3 * k + c] != 0.f) atomicAdd(dL_dtetpoints + 3 * vid[k] + c, v[3 * k + c]);
        }
    }
}

__global__ void __launch_bounds__(GEO_THREADS, DEFORM_BWD_MIN_CTAS) deform_backward_kernel(
    int N, int T, const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
    const int64_t *__restrict__ tetra_id, const float *__restrict__ barys, const float *__restrict__ cgrad,
    const float *__restrict__ scales, const float *__restrict__ rotation, const float *__restrict__ dL_dmeans,
    const float *__restrict__ dL_dcov, float *__restrict__ dL_dtetpoints, float *__restrict__ dL_dbarys,
    float *__restrict__ dL_dscales, float *__restrict__ dL_drotation, const int32_t *__restrict__ order,
    const DeformExtra ex) {
    if ((int)blockIdx.x >= ex.gaussian_blocks) {   // appended CTAs: gradient of the cage energy into the cage vertices
        const int tt = ((int)blockIdx.x - ex.gaussian_blocks) * GEO_THREADS + (int)threadIdx.x;
        if (tt < T && dL_dtetpoints) fem::fem_backward_tet(tt, tetpoints, tets, ex.dn_inv, ex.dL_denergy, dL_dtetpoints, ex.grad_pitch);
        return;
    }
    const int i = blockIdx.x * GEO_THREADS + threadIdx.x;
    // the sparse form: `order` lists the *count Gaussians whose incoming gradient is not zero (deform_select_kernel)
    const int n_rows = ex.count ? (int)__ldg(ex.count) : N;
    if ((int)(blockIdx.x * GEO_THREADS) >= n_rows) return;
    const bool live = i < n_rows;
    // With `order` (a permutation that sorts tetra_id) neighbouring lanes share tets: their cage-vertex gradients are
    // summed inside the warp and one lane per tet issues the atomics (~N/T times fewer L2 atomics).
    const int n = live ? (order ? order[i] : i) : 0;
    int64_t t = -1 - (int64_t)(threadIdx.x & 31);   // dead lanes: distinct keys, never merged with anything
    int64_t vid[4] = {0, 0, 0, 0};
    float dvf[12];
#pragma unroll
    for (int k = 0; k < 12; k++) dvf[k] = 0.f;

    // The [N,K] streams.  In natural order the warp's rows are contiguous and go through shared memory as coalesced
    // requests (every lane takes part); with `order` the rows are gathered, one lane one row.
    __shared__ float s_stage[GEO_THREADS / 32][STAGE_FLOATS];
    const uint32_t lane = threadIdx.x & 31;
    float *s_w = s_stage[threadIdx.x >> 5];
    const size_t row0 = (size_t)blockIdx.x * GEO_THREADS + (threadIdx.x & ~31u);
    float cg9[9], sraw[3], dm[3], g6[6], ds_out[3] = {0.f, 0.f, 0.f};
    if (!order) {
        float r9[9], r6[6], rm[3], r3[3], rd[3] = {0.f, 0.f, 0.f};
        warp_rows_fetch<9>(cgrad, row0, (size_t)N, lane, r9);
        warp_rows_fetch<6>(dL_dcov, row0, (size_t)N, lane, r6);
        warp_rows_fetch<3>(dL_dmeans, row0, (size_t)N, lane, rm);
        warp_rows_fetch<3>(scales, row0, (size_t)N, lane, r3);
        if (ex.delta_scale) warp_rows_fetch<3>(ex.delta_scale, row0, (size_t)N, lane, rd);
        if (live) t = __ldg(tetra_id + n);
        r3[0] += rd[0]; r3[1] += rd[1]; r3[2] += rd[2];
        warp_rows_take<9>(s_w, lane, r9, cg9);
        warp_rows_take<6>(s_w, lane, r6, g6);
        warp_rows_take<3>(s_w, lane, rm, dm);
        warp_rows_take<3>(s_w, lane, r3, sraw);
    } else if (live) {
        t = __ldg(tetra_id + n);
#pragma unroll
        for (int k = 0; k < 9; k++) cg9[k] = cgrad[9 * (size_t)n + k];
#pragma unroll
        for (int k = 0; k < 6; k++) g6[k] = dL_dcov[6 * (size_t)n + k];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            dm[k] = dL_dmeans[3 * (size_t)n + k];
            sraw[k] = scales[3 * (size_t)n + k];
            if (ex.delta_scale) sraw[k] += ex.delta_scale[3 * (size_t)n + k];
        }
    }

    if (live) {
        if (t < 0) t += T;
        Tet4 X;
        load_tet(tetpoints, tets, t, vid, X);
        float bw[4], s[3];
        float4 q;
        load_gaussian(ex, barys, sraw, rotation, (size_t)n, bw, s, q);

        // recompute the forward
        float G[3][3], Cg[3][3], J[3][3];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            G[r][0] = X.v[3][r] - X.v[0][r]; G[r][1] = X.v[2][r] - X.v[0][r]; G[r][2] = X.v[1][r] - X.v[0][r];
        }
#pragma unroll
        for (int k = 0; k < 9; k++) Cg[k / 3][k % 3] = cg9[k];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int j = 0; j < 3; j++) J[r][j] = G[r][0] * Cg[0][j] + G[r][1] * Cg[1][j] + G[r][2] * Cg[2][j];
        const float inv = 1.f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        const float qr = q.x * inv, qx = q.y * inv, qy = q.z * inv, qz = q.w * inv;
        float Rm[3][3];
        rotation_from_unit_quat(qr, qx, qy, qz, Rm);
        float JR[3][3], B[3][3];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int k = 0; k < 3; k++) {
                JR[r][k] = J[r][0] * Rm[0][k] + J[r][1] * Rm[1][k] + J[r][2] * Rm[2][k];
                B[r][k] = JR[r][k] * s[k];
            }
        // cov = B B^T, only the upper triangle is an output: dB = (D + D^T) B with D the upper-triangular gradient
        const float Ds[3][3] = {{2.f * g6[0], g6[1], g6[2]}, {g6[1], 2.f * g6[3], g6[4]}, {g6[2], g6[4], 2.f * g6[5]}};
        float dB[3][3];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int k = 0; k < 3; k++) dB[r][k] = Ds[r][0] * B[0][k] + Ds[r][1] * B[1][k] + Ds[r][2] * B[2][k];
        // B = (J R) diag(s)
        float ds[3], dJR[3][3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            ds[k] = dB[0][k] * JR[0][k] + dB[1][k] * JR[1][k] + dB[2][k] * JR[2][k];
#pragma unroll
            for (int r = 0; r < 3; r++) dJR[r][k] = dB[r][k] * s[k];
        }
        // JR = J R : dJ = dJR R^T, dR = J^T dJR
        float dJ[3][3], A[3][3];
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int j = 0; j < 3; j++) {
                dJ[r][j] = dJR[r][0] * Rm[j][0] + dJR[r][1] * Rm[j][1] + dJR[r][2] * Rm[j][2];
                A[r][j] = J[0][r] * dJR[0][j] + J[1][r] * dJR[1][j] + J[2][r] * dJR[2][j];
            }
        // through exp when the scales came in pre-activation: d exp(x) = exp(x) dx
#pragma unroll
        for (int k = 0; k < 3; k++) ds_out[k] = ex.log_scales ? ds[k] * s[k] : ds[k];
        if (dL_dscales && order) {
#pragma unroll
            for (int k = 0; k < 3; k++) dL_dscales[3 * (size_t)n + k] = ds_out[k];
        }
        if (dL_drotation) {
            const float r = qr, x = qx, y = qy, z = qz;
            float dq[4];
            dq[0] = 2.f * (-z * A[0][1] + y * A[0][2] + z * A[1][0] - x * A[1][2] - y * A[2][0] + x * A[2][1]);
            dq[1] = 2.f * (y * A[0][1] + z * A[0][2] + y * A[1][0] - 2.f * x * A[1][1] - r * A[1][2] + z * A[2][0] + r * A[2][1] - 2.f * x * A[2][2]);
            dq[2] = 2.f * (-2.f * y * A[0][0] + x * A[0][1] + r * A[0][2] + x * A[1][0] + z * A[1][2] - r * A[2][0] + z * A[2][1] - 2.f * y * A[2][2]);
            dq[3] = 2.f * (-2.f * z * A[0][0] - r * A[0][1] + x * A[0][2] + r * A[1][0] - 2.f * z * A[1][1] + y * A[1][2] + x * A[2][0] + y * A[2][1]);
            // through q / |q|
            const float dotp = r * dq[0] + x * dq[1] + y * dq[2] + z * dq[3];
            *reinterpret_cast<float4 *>(dL_drotation + 4 * (size_t)n) =
                make_float4((dq[0] - r * dotp) * inv, (dq[1] - x * dotp) * inv, (dq[2] - y * dotp) * inv, (dq[3] - z * dotp) * inv);
        }
        if (dL_dbarys) {
            float db[4];
#pragma unroll
            for (int k = 0; k < 4; k++) db[k] = X.v[k][0] * dm[0] + X.v[k][1] * dm[1] + X.v[k][2] * dm[2];
            *reinterpret_cast<float4 *>(dL_dbarys + 4 * (size_t)n) = make_float4(db[0], db[1], db[2], db[3]);
        }
        if (dL_dtetpoints) {
            // J = G Cg : dG = dJ Cg^T ; columns of G are (v3-v0, v2-v0, v1-v0)
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const float g0 = dJ[r][0] * Cg[0][0] + dJ[r][1] * Cg[0][1] + dJ[r][2] * Cg[0][2];
                const float g1 = dJ[r][0] * Cg[1][0] + dJ[r][1] * Cg[1][1] + dJ[r][2] * Cg[1][2];
                const float g2 = dJ[r][0] * Cg[2][0] + dJ[r][1] * Cg[2][1] + dJ[r][2] * Cg[2][2];
                dvf[9 + r] = g0 + bw[3] * dm[r];
                dvf[6 + r] = g1 + bw[2] * dm[r];
                dvf[3 + r] = g2 + bw[1] * dm[r];
                dvf[0 + r] = -(g0 + g1 + g2) + bw[0] * dm[r];
            }
        }
    }
    if (dL_dscales && !order) warp_rows_store<3>(dL_dscales, row0, (size_t)N, s_w, lane, ds_out);
    if (!dL_dtetpoints) return;
    if ((order && !ex.plain_list) || ex.tet_sorted) {
        segmented_vertex_scatter(t, vid, dvf, dL_dtetpoints, ex.grad_pitch);   // every lane, one call site
    } else if (live) {
        if (ex.grad_pitch == 4) {
            // padded rows: a vertex's three components go out as ONE 16-byte vector reduction (4 per Gaussian, not 12)
#pragma unroll
            for (int k = 0; k < 4; k++) red_add_v4(dL_dtetpoints + 4 * vid[k], dvf[3 * k], dvf[3 * k + 1], dvf[3 * k + 2], 0.f);
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++)
#pragma unroll
                for (int c = 0; c < 3; c++) atomicAdd(dL_dtetpoints + 3 * vid[k] + c, dvf[3 * k + c]);
        }
    }
}

// The sparse form of the backward, step 1.  Most Gaussians of a rendered view come back with an exactly-zero gradient:
// whatever lies behind the surface that saturated the pixels was never blended (measured on the 1 M-Gaussian avatar:
// 86-93 % per camera).  This kernel streams the incoming (dL/dmean, dL/dcov) once, writes the zero gradients of those
// Gaussians itself and lists the others (one append per CTA; a CTA's batch stays in row order, so tet-sorted rows
// still meet their neighbours in the warp sums); deform_backward_kernel then runs over the list alone.
__global__ void __launch_bounds__(GEO_THREADS) deform_select_kernel(int N, const float *__restrict__ dL_dmeans,
                                                                    const float *__restrict__ dL_dcov,
                                                                    float *__restrict__ dL_dbarys, float *__restrict__ dL_dscales,
                                                                    float *__restrict__ dL_drotation, int32_t *__restrict__ list,
                                                                    uint32_t *__restrict__ count) {
    __shared__ float s_stage[GEO_THREADS / 32][32 * 6];
    const uint32_t lane = threadIdx.x & 31;
    float *s_w = s_stage[threadIdx.x >> 5];
    const size_t row0 = (size_t)blockIdx.x * GEO_THREADS + (threadIdx.x & ~31u);
    const size_t i = row0 + lane;
    const bool live = i < (size_t)N;
    float r6[6], rm[3], z6[6], z3[3];
    warp_rows_fetch<6>(dL_dcov, row0, (size_t)N, lane, r6);
    warp_rows_fetch<3>(dL_dmeans, row0, (size_t)N, lane, rm);
    warp_rows_take<6>(s_w, lane, r6, z6);
    warp_rows_take<3>(s_w, lane, rm, z3);
    const bool nz = live && (z3[0] != 0.f || z3[1] != 0.f || z3[2] != 0.f || z6[0] != 0.f || z6[1] != 0.f || z6[2] != 0.f ||
                             z6[3] != 0.f || z6[4] != 0.f || z6[5] != 0.f);
    if (live && !nz) {
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        if (dL_dbarys) *reinterpret_cast<float4 *>(dL_dbarys + 4 * i) = zero;
        if (dL_drotation) *reinterpret_cast<float4 *>(dL_drotation + 4 * i) = zero;
        if (dL_dscales) { dL_dscales[3 * i] = 0.f; dL_dscales[3 * i + 1] = 0.f; dL_dscales[3 * i + 2] = 0.f; }
    }
    // one append per CTA (a counter that every warp of the grid adds to serialises in the L2: 35 us for 31 k adds)
    __shared__ uint32_t s_cnt[GEO_THREADS / 32], s_base;
    const uint32_t mask = __ballot_sync(0xffffffffu, nz), warp = threadIdx.x >> 5;
    if (lane == 0) s_cnt[warp] = (uint32_t)__popc(mask);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < GEO_THREADS / 32; w++) total += s_cnt[w];
        s_base = total ? atomicAdd(count, total) : 0u;
    }
    __syncthreads();
    if (!nz) return;
    uint32_t off = s_base;
    for (uint32_t w = 0; w < warp; w++) off += s_cnt[w];
    list[off + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = (int32_t)i;
}

}  // namespace
}  // namespace d3ga

using namespace d3ga;

static int deform_forward_impl(int N, int V, int T, const float *tetpoints, const int64_t *tets, const int64_t *tetra_id,
                               const float *barys, const float *canonical_gradient, const float *scales,
                               const float *rotation, float *means3D, float *cov3D, DeformExtra ex, d3ga_stream_t stream) {
    if (N < 0 || V <= 0 || T <= 0) return D3GA_EINVAL;
    if (!tetpoints || !tets) return D3GA_EINVAL;
    if (N > 0 && (!tetra_id || !barys || !canonical_gradient || !scales || !rotation || !means3D || !cov3D)) return D3GA_EINVAL;
    if ((ex.dn_inv != nullptr) != (ex.energy != nullptr)) return D3GA_EINVAL;
    ex.gaussian_blocks = (N + GEO_THREADS - 1) / GEO_THREADS;
    const int blocks = ex.gaussian_blocks + (ex.dn_inv ? (T + GEO_THREADS - 1) / GEO_THREADS : 0);
    if (blocks == 0) return D3GA_OK;
    ProfScope prof(K_DEFORM_FWD, (cudaStream_t)stream);
    deform_forward_kernel<<<blocks, GEO_THREADS, 0, (cudaStream_t)stream>>>(
        N, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, means3D, cov3D, ex);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_deform_forward(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                   const int64_t *tetra_id, const float *barys, const float *canonical_gradient,
                                   const float *scales, const float *rotation, float *means3D, float *cov3D,
                                   d3ga_stream_t stream) {
    if (N == 0) return (N < 0 || V <= 0 || T <= 0) ? D3GA_EINVAL : D3GA_OK;
    return deform_forward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, means3D,
                               cov3D, DeformExtra{}, stream);
}

extern "C" int d3ga_deform_forward_fused(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                         const int64_t *tetra_id, const float *barys, const float *delta_bary,
                                         const float *canonical_gradient, const float *scales, const float *delta_scale,
                                         int log_scales, const float *rotation, const float *delta_rot,
                                         const float *dn_inv, float *energy, float *means3D, float *cov3D,
                                         d3ga_stream_t stream) {
    DeformExtra ex{};
    ex.delta_bary = delta_bary; ex.delta_scale = delta_scale; ex.delta_rot = delta_rot; ex.log_scales = log_scales ? 1 : 0;
    ex.dn_inv = dn_inv; ex.energy = energy;
    return deform_forward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, means3D,
                               cov3D, ex, stream);
}

static int deform_backward_impl(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                const int64_t *tetra_id, const float *barys, const float *canonical_gradient,
                                const float *scales, const float *rotation, const float *dL_dmeans3D,
                                const float *dL_dcov3D, float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales,
                                float *dL_drotation, const int32_t *order, DeformExtra ex, d3ga_stream_t stream,
                                void *sparse_scratch = nullptr) {
    if (N < 0 || V <= 0 || T <= 0) return D3GA_EINVAL;
    if (!tetpoints || !tets) return D3GA_EINVAL;
    if (N > 0 && (!tetra_id || !barys || !canonical_gradient || !scales || !rotation || !dL_dmeans3D || !dL_dcov3D))
        return D3GA_EINVAL;
    if (ex.dL_denergy && !ex.dn_inv) return D3GA_EINVAL;
    if (ex.grad_pitch != 4) ex.grad_pitch = 3;
    ex.gaussian_blocks = (N + GEO_THREADS - 1) / GEO_THREADS;
    const int blocks = ex.gaussian_blocks + (ex.dL_denergy ? (T + GEO_THREADS - 1) / GEO_THREADS : 0);
    if (blocks == 0) return D3GA_OK;
    if (sparse_scratch && N > 0) {
        if (order) return D3GA_EINVAL;
        uint32_t *count = reinterpret_cast<uint32_t *>(sparse_scratch);
        int32_t *list = reinterpret_cast<int32_t *>(reinterpret_cast<uint8_t *>(sparse_scratch) + 256);
        D3GA_CUDA(cudaMemsetAsync(count, 0, 4, (cudaStream_t)stream));
        {
            ProfScope prof(K_DEFORM_SELECT, (cudaStream_t)stream);
            deform_select_kernel<<<ex.gaussian_blocks, GEO_THREADS, 0, (cudaStream_t)stream>>>(
                N, dL_dmeans3D, dL_dcov3D, dL_dbarys, dL_dscales, dL_drotation, list, count);
        }
        D3GA_LAUNCH_CHECK();
        order = list;
        ex.count = count;
        ex.plain_list = ex.tet_sorted ? 0 : 1;     // rows in tet order: a warp's neighbours still share tets
        ex.tet_sorted = 0;
    }
    ProfScope prof(K_DEFORM_BWD, (cudaStream_t)stream);
    deform_backward_kernel<<<blocks, GEO_THREADS, 0, (cudaStream_t)stream>>>(
        N, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation, dL_dmeans3D, dL_dcov3D,
        dL_dtetpoints, dL_dbarys, dL_dscales, dL_drotation, order, ex);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_deform_backward(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                    const int64_t *tetra_id, const float *barys, const float *canonical_gradient,
                                    const float *scales, const float *rotation, const float *dL_dmeans3D,
                                    const float *dL_dcov3D, float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales,
                                    float *dL_drotation, d3ga_stream_t stream) {
    if (N == 0) return (N < 0 || V <= 0 || T <= 0) ? D3GA_EINVAL : D3GA_OK;
    return deform_backward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation,
                                dL_dmeans3D, dL_dcov3D, dL_dtetpoints, dL_dbarys, dL_dscales, dL_drotation, nullptr,
                                DeformExtra{}, stream);
}

extern "C" int d3ga_deform_backward_ordered(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                            const int64_t *tetra_id, const float *barys,
                                            const float *canonical_gradient, const float *scales,
                                            const float *rotation, const float *dL_dmeans3D, const float *dL_dcov3D,
                                            float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales,
                                            float *dL_drotation, const int32_t *order, d3ga_stream_t stream) {
    if (!order) return D3GA_EINVAL;
    if (N == 0) return (N < 0 || V <= 0 || T <= 0) ? D3GA_EINVAL : D3GA_OK;
    return deform_backward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation,
                                dL_dmeans3D, dL_dcov3D, dL_dtetpoints, dL_dbarys, dL_dscales, dL_drotation, order,
                                DeformExtra{}, stream);
}

extern "C" int d3ga_deform_backward_fused(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                          const int64_t *tetra_id, const float *barys, const float *delta_bary,
                                          const float *canonical_gradient, const float *scales, const float *delta_scale,
                                          int log_scales, const float *rotation, const float *delta_rot,
                                          const float *dn_inv, const float *dL_dmeans3D, const float *dL_dcov3D,
                                          const float *dL_denergy, float *dL_dtetpoints, float *dL_dbarys,
                                          float *dL_dscales, float *dL_drotation, const int32_t *order,
                                          int rows_tet_sorted, int dL_dtetpoints_pitch, d3ga_stream_t stream) {
    if (dL_dtetpoints_pitch != 3 && dL_dtetpoints_pitch != 4) return D3GA_EINVAL;
    if (order && rows_tet_sorted) return D3GA_EINVAL;
    DeformExtra ex{};
    ex.grad_pitch = dL_dtetpoints_pitch;
    ex.tet_sorted = rows_tet_sorted ? 1 : 0;
    ex.delta_bary = delta_bary; ex.delta_scale = delta_scale; ex.delta_rot = delta_rot; ex.log_scales = log_scales ? 1 : 0;
    ex.dn_inv = dn_inv; ex.dL_denergy = dL_denergy;
    return deform_backward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation,
                                dL_dmeans3D, dL_dcov3D, dL_dtetpoints, dL_dbarys, dL_dscales, dL_drotation, order, ex,
                                stream);
}

extern "C" uint64_t d3ga_deform_backward_scratch_bytes(int N) { return 256 + 4ull * (uint64_t)(N > 0 ? N : 1); }

extern "C" int d3ga_deform_backward_fused_sparse(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                                 const int64_t *tetra_id, const float *barys, const float *delta_bary,
                                                 const float *canonical_gradient, const float *scales,
                                                 const float *delta_scale, int log_scales, const float *rotation,
                                                 const float *delta_rot, const float *dn_inv, const float *dL_dmeans3D,
                                                 const float *dL_dcov3D, const float *dL_denergy, float *dL_dtetpoints,
                                                 float *dL_dbarys, float *dL_dscales, float *dL_drotation,
                                                 int rows_tet_sorted, int dL_dtetpoints_pitch, void *scratch,
                                                 d3ga_stream_t stream) {
    if (dL_dtetpoints_pitch != 3 && dL_dtetpoints_pitch != 4) return D3GA_EINVAL;
    if (!scratch) return D3GA_EINVAL;
    DeformExtra ex{};
    ex.grad_pitch = dL_dtetpoints_pitch;
    ex.tet_sorted = rows_tet_sorted ? 1 : 0;
    ex.delta_bary = delta_bary; ex.delta_scale = delta_scale; ex.delta_rot = delta_rot; ex.log_scales = log_scales ? 1 : 0;
    ex.dn_inv = dn_inv; ex.dL_denergy = dL_denergy;
    return deform_backward_impl(N, V, T, tetpoints, tets, tetra_id, barys, canonical_gradient, scales, rotation,
                                dL_dmeans3D, dL_dcov3D, dL_dtetpoints, dL_dbarys, dL_dscales, dL_drotation, nullptr, ex,
                                stream, scratch);
}
