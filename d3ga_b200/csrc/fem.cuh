// fem.cuh -- per-tet device code of the stable neo-Hookean cage energy (lib/cage.py:349-361), shared by the stand-alone
// kernels of fem.cu and by the deformation kernels of deform.cu, which run it as extra CTAs of their own launch (it is
// the other consumer of the gather tetpoints[tets] and the other source of gradient into tetpoints: SURVEY.md 8(f)-2).
#pragma once
#include "common.cuh"

namespace d3ga {
namespace fem {


__device__ __forceinline__ void load_F(const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
                                       const float *__restrict__ dn_inv, int t, int64_t (&vid)[4], float (&Dn)[3][3],
                                       float (&F)[3][3]) {
    float v[4][3];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        vid[k] = tets[4 * (size_t)t + k];
#pragma unroll
        for (int c = 0; c < 3; c++) v[k][c] = __ldg(tetpoints + 3 * vid[k] + c);
    }
    float Ds[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++) { Ds[r][0] = v[3][r] - v[0][r]; Ds[r][1] = v[2][r] - v[0][r]; Ds[r][2] = v[1][r] - v[0][r]; }
#pragma unroll
    for (int k = 0; k < 9; k++) Dn[k / 3][k % 3] = dn_inv[9 * (size_t)t + k];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) F[r][c] = Ds[r][0] * Dn[0][c] + Ds[r][1] * Dn[1][c] + Ds[r][2] * Dn[2][c];
}

__device__ __forceinline__ float det3(const float (&F)[3][3]) {
    return F[0][0] * (F[1][1] * F[2][2] - F[1][2] * F[2][1]) - F[0][1] * (F[1][0] * F[2][2] - F[1][2] * F[2][0]) +
           F[0][2] * (F[1][0] * F[2][1] - F[1][1] * F[2][0]);
}

__device__ __forceinline__ void fem_forward_tet(int t, const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
                                                const float *__restrict__ dn_inv, float *__restrict__ energy) {
    int64_t vid[4];
    float Dn[3][3], F[3][3];
    load_F(tetpoints, tets, dn_inv, t, vid, Dn, F);
    const float J = det3(F);
    float tr = 0.f;
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) tr += F[r][c] * F[r][c];
    energy[t] = 0.5f * (J - 1.f) * (J - 1.f) + 0.5f * (tr - 3.f);
}

__device__ __forceinline__ void fem_backward_tet(int t, const float *__restrict__ tetpoints, const int64_t *__restrict__ tets,
                                                 const float *__restrict__ dn_inv, const float *__restrict__ dL_denergy,
                                                 float *__restrict__ dL_dtetpoints, int pitch = 3) {
    const float g = dL_denergy[t];
    if (g == 0.f) return;
    int64_t vid[4];
    float Dn[3][3], F[3][3];
    load_F(tetpoints, tets, dn_inv, t, vid, Dn, F);
    const float J = det3(F);
    // dE/dF = (J - 1) cof(F) + F
    float C[3][3];
    C[0][0] = F[1][1] * F[2][2] - F[1][2] * F[2][1]; C[0][1] = F[1][2] * F[2][0] - F[1][0] * F[2][2]; C[0][2] = F[1][0] * F[2][1] - F[1][1] * F[2][0];
    C[1][0] = F[0][2] * F[2][1] - F[0][1] * F[2][2]; C[1][1] = F[0][0] * F[2][2] - F[0][2] * F[2][0]; C[1][2] = F[0][1] * F[2][0] - F[0][0] * F[2][1];
    C[2][0] = F[0][1] * F[1][2] - F[0][2] * F[1][1]; C[2][1] = F[0][2] * F[1][0] - F[0][0] * F[1][2]; C[2][2] = F[0][0] * F[1][1] - F[0][1] * F[1][0];
    float dF[3][3];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int c = 0; c < 3; c++) dF[r][c] = g * ((J - 1.f) * C[r][c] + F[r][c]);
    // F = Ds Dn : dDs = dF Dn^T ; columns of Ds are (v3-v0, v2-v0, v1-v0)
    float dv[4][3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const float d0 = dF[r][0] * Dn[0][0] + dF[r][1] * Dn[0][1] + dF[r][2] * Dn[0][2];
        const float d1 = dF[r][0] * Dn[1][0] + dF[r][1] * Dn[1][1] + dF[r][2] * Dn[1][2];
        const float d2 = dF[r][0] * Dn[2][0] + dF[r][1] * Dn[2][1] + dF[r][2] * Dn[2][2];
        dv[3][r] = d0; dv[2][r] = d1; dv[1][r] = d2; dv[0][r] = -(d0 + d1 + d2);
    }
    if (pitch == 4) {   // padded accumulator: a vertex's three components as one 16-byte vector reduction
#pragma unroll
        for (int k = 0; k < 4; k++) red_add_v4(dL_dtetpoints + 4 * vid[k], dv[k][0], dv[k][1], dv[k][2], 0.f);
    } else {
#pragma unroll
        for (int r = 0; r < 3; r++) {
            atomicAdd(dL_dtetpoints + 3 * vid[3] + r, dv[3][r]);
            atomicAdd(dL_dtetpoints + 3 * vid[2] + r, dv[2][r]);
            atomicAdd(dL_dtetpoints + 3 * vid[1] + r, dv[1][r]);
            atomicAdd(dL_dtetpoints + 3 * vid[0] + r, dv[0][r]);
        }
    }
}

}  // namespace fem
}  // namespace d3ga
