// fem.cu -- SURVEY.md section 8(f) rank 2: the stable neo-Hookean cage energy and its gradient, one kernel each way.
//
// Replaces CageBase.fem_energy (lib/cage.py:349-361) -- Ds = G(points[tets]); F = Ds @ Dn_inv;
// E = 0.5 (det F - 1)^2 + 0.5 (tr(F^T F) - 3) per tet -- and its autograd backward (about ten torch kernels forward
// and twice that backward over [T,3,3] temporaries, called every frame at models/cage_net.py:225).  It is the other
// consumer of the cage-vertex gather tetpoints[tets] and the other source of gradient into tetpoints.
// One thread per tet.  G uses the edge order of tetra_sampler/tetra.py:74-80: columns (v3-v0, v2-v0, v1-v0).
// Algorithmic bytes per tet: 32 (ids) + 48 (vertices, L2) + 36 (Dn_inv) read, 4 written forward; + 48 B scatter backward.
#include "fem.cuh"

namespace d3ga {
namespace {

__global__ void __launch_bounds__(GEO_THREADS) fem_energy_forward_kernel(int T, const float *__restrict__ tetpoints,
                                                                        const int64_t *__restrict__ tets,
                                                                        const float *__restrict__ dn_inv,
                                                                        float *__restrict__ energy) {
    const int t = blockIdx.x * GEO_THREADS + threadIdx.x;
    if (t < T) fem::fem_forward_tet(t, tetpoints, tets, dn_inv, energy);
}

__global__ void __launch_bounds__(GEO_THREADS) fem_energy_backward_kernel(int T, const float *__restrict__ tetpoints,
                                                                         const int64_t *__restrict__ tets,
                                                                         const float *__restrict__ dn_inv,
                                                                         const float *__restrict__ dL_denergy,
                                                                         float *__restrict__ dL_dtetpoints) {
    const int t = blockIdx.x * GEO_THREADS + threadIdx.x;
    if (t < T) fem::fem_backward_tet(t, tetpoints, tets, dn_inv, dL_denergy, dL_dtetpoints);
}

}  // namespace
}  // namespace d3ga

using namespace d3ga;

extern "C" int d3ga_fem_energy_forward(int T, int V, const float *tetpoints, const int64_t *tets, const float *dn_inv,
                                       float *energy, d3ga_stream_t stream) {
    if (T < 0 || V <= 0) return D3GA_EINVAL;
    if (T == 0) return D3GA_OK;
    if (!tetpoints || !tets || !dn_inv || !energy) return D3GA_EINVAL;
    ProfScope prof(K_FEM_FWD, (cudaStream_t)stream);
    fem_energy_forward_kernel<<<(T + GEO_THREADS - 1) / GEO_THREADS, GEO_THREADS, 0, (cudaStream_t)stream>>>(
        T, tetpoints, tets, dn_inv, energy);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_fem_energy_backward(int T, int V, const float *tetpoints, const int64_t *tets, const float *dn_inv,
                                        const float *dL_denergy, float *dL_dtetpoints, d3ga_stream_t stream) {
    if (T < 0 || V <= 0) return D3GA_EINVAL;
    if (T == 0) return D3GA_OK;
    if (!tetpoints || !tets || !dn_inv || !dL_denergy || !dL_dtetpoints) return D3GA_EINVAL;
    ProfScope prof(K_FEM_BWD, (cudaStream_t)stream);
    fem_energy_backward_kernel<<<(T + GEO_THREADS - 1) / GEO_THREADS, GEO_THREADS, 0, (cudaStream_t)stream>>>(
        T, tetpoints, tets, dn_inv, dL_denergy, dL_dtetpoints);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}
