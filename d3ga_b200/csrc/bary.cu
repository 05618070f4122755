// bary.cu -- row A0 of SURVEY.md section 8: embed points in the tets of a cage (init-time, once per garment).
//
// Replaces tetra_sampler.compute_bary (submodules/tetra-sampler/tetra_sampler/barys.py:9-33), i.e.
//   (1) bvh_distance_queries.BVH()(triangles, points) -> nearest unique tet face per point
//       (submodules/bvh/src/bvh_cuda_op.cu:1065-1269: LBVH build + stack traversal, a device sync after every
//       launch and five cudaMallocs per call), and
//   (2) barycentric_kernel (src/sampler_points_kernel.cu:279-313): test the <= 2 tets adjacent to that face with
//       point_in_tet / same_side (include/utils.h:394-423), first hit wins (get_id, utils.h:468-481), then
//       barycentric() (utils.h:425-451) -- `1.0 / triple` is a DOUBLE divide rounded to float, kept here.
// Step (1) is an exact brute-force search tiled through shared memory: for the cage sizes D3GA uses (10^4..10^5
// faces, 10^4..10^5 points) that is a few ms on a B200 and needs no tree, no sort, no allocation and no host sync.
// Ties between equidistant faces go to the lowest face index (the reference's BVH resolves them by traversal
// order); either way the containing tet is adjacent to the chosen face, so (tetra_id, barys) agree.
#include "common.cuh"

namespace d3ga {

namespace {

constexpr int BARY_THREADS = 256;
constexpr int FACE_TILE = 256;

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

// squared distance from p to triangle (a,b,c): closest-point-by-Voronoi-region, the form the reference's BVH query
// evaluates (submodules/bvh/src/bvh_cuda_op.cu:209-281)
__device__ float point_triangle_dist2(V3 p, V3 a, V3 b, V3 c) {
    const V3 ab = b - a, ac = c - a, ap = p - a;
    const float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) return dot(ap, ap);
    const V3 bp = p - b;
    const float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) return dot(bp, bp);
    const float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) {
        const float v = d1 / (d1 - d3);
        const V3 d = p - (a + v * ab);
        return dot(d, d);
    }
    const V3 cp = p - c;
    const float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) return dot(cp, cp);
    const float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) {
        const float w = d2 / (d2 - d6);
        const V3 d = p - (a + w * ac);
        return dot(d, d);
    }
    const float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        const float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        const V3 d = p - (b + w * (c - b));
        return dot(d, d);
    }
    const float denom = 1.f / (va + vb + vc);
    const float v = vb * denom, w = vc * denom;
    const V3 d = p - (a + v * ab + w * ac);
    return dot(d, d);
}

__device__ __forceinline__ bool same_side(V3 v1, V3 v2, V3 v3, V3 v4, V3 p) {
    const V3 normal = cross(v2 - v1, v3 - v1);
    return signbit(dot(normal, v4 - v1)) == signbit(dot(normal, p - v1));
}
__device__ __forceinline__ bool point_in_tet(const V3 (&v)[4], V3 p) {
    return same_side(v[0], v[1], v[2], v[3], p) && same_side(v[1], v[2], v[3], v[0], p) &&
           same_side(v[2], v[3], v[0], v[1], p) && same_side(v[3], v[0], v[1], v[2], p);
}
__device__ __forceinline__ void load_tet(const float *tetras, int id, V3 (&v)[4]) {
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = {tetras[12 * (size_t)id + 3 * k], tetras[12 * (size_t)id + 3 * k + 1], tetras[12 * (size_t)id + 3 * k + 2]};
}

__global__ void __launch_bounds__(BARY_THREADS) compute_bary_kernel(
    int N, int F, const float *__restrict__ points, const float *__restrict__ tetras,
    const float *__restrict__ triangles, const int2 *__restrict__ tri_to_tetra, float4 *__restrict__ barys,
    int32_t *__restrict__ tetra_id, int32_t *__restrict__ active, int32_t *__restrict__ closest_face) {
    __shared__ float s_tri[FACE_TILE * 9];
    const int n = blockIdx.x * BARY_THREADS + threadIdx.x;
    const bool ok = n < N;
    V3 p = {0.f, 0.f, 0.f};
    if (ok) p = {points[3 * (size_t)n], points[3 * (size_t)n + 1], points[3 * (size_t)n + 2]};
    float best = INFINITY;
    int best_face = -1;
    for (int f0 = 0; f0 < F; f0 += FACE_TILE) {
        const int nf = min(FACE_TILE, F - f0);
        __syncthreads();
        for (int k = threadIdx.x; k < nf * 9; k += BARY_THREADS) s_tri[k] = triangles[(size_t)f0 * 9 + k];
        __syncthreads();
        if (ok) {
            for (int f = 0; f < nf; f++) {
                const float *t = s_tri + 9 * f;
                const float d = point_triangle_dist2(p, {t[0], t[1], t[2]}, {t[3], t[4], t[5]}, {t[6], t[7], t[8]});
                if (d < best) { best = d; best_face = f0 + f; }
            }
        }
    }
    if (!ok) return;
    if (closest_face) closest_face[n] = best_face;
    int id = -1;
    if (best_face >= 0) {
        const int2 nb = tri_to_tetra[best_face];
        V3 v[4];
        if (nb.x != -1) { load_tet(tetras, nb.x, v); if (point_in_tet(v, p)) id = nb.x; }
        if (id == -1 && nb.y != -1) { load_tet(tetras, nb.y, v); if (point_in_tet(v, p)) id = nb.y; }
    }
    if (id == -1) {
        barys[n] = make_float4(0.f, 0.f, 0.f, 0.f);
        active[n] = -1;
        tetra_id[n] = -1;
        return;
    }
    V3 v[4];
    load_tet(tetras, id, v);
    const V3 a = v[0], b = v[1], c = v[2], d = v[3];
    const V3 vap = p - a, vbp = p - b, vab = b - a, vac = c - a, vad = d - a, vbc = c - b, vbd = d - b;
    const float va6 = dot(vbp, cross(vbd, vbc)), vb6 = dot(vap, cross(vac, vad));
    const float vc6 = dot(vap, cross(vad, vab)), vd6 = dot(vap, cross(vab, vac));
    const float v6 = (float)(1.0 / (double)dot(vab, cross(vac, vad)));
    barys[n] = make_float4(va6 * v6, vb6 * v6, vc6 * v6, vd6 * v6);
    tetra_id[n] = id;
    active[n] = 1;
}

}  // namespace
}  // namespace d3ga

using namespace d3ga;

extern "C" int d3ga_compute_bary(int N, int T, int F, const float *points, const float *tetras, const float *triangles,
                                 const int32_t *tri_to_tetra, float *barys, int32_t *tetra_id, int32_t *active,
                                 int32_t *closest_face, d3ga_stream_t stream) {
    if (N < 0 || T <= 0 || F <= 0) return D3GA_EINVAL;
    if (!points || !tetras || !triangles || !tri_to_tetra || !barys || !tetra_id || !active) return D3GA_EINVAL;
    if (N == 0) return D3GA_OK;
    ProfScope prof(K_BARY, (cudaStream_t)stream);
    compute_bary_kernel<<<(N + BARY_THREADS - 1) / BARY_THREADS, BARY_THREADS, 0, (cudaStream_t)stream>>>(
        N, F, points, tetras, triangles, reinterpret_cast<const int2 *>(tri_to_tetra),
        reinterpret_cast<float4 *>(barys), tetra_id, active, closest_face);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}
