/*
 * oracle/tet_oracle.c -- TEST INFRASTRUCTURE ONLY.
 * CPU restatement of tetra_sampler.compute_bary (submodules/tetra-sampler/tetra_sampler/barys.py:9-33):
 *   nearest unique tet face per point  -- the query bvh_distance_queries.BVH answers (bvh_cuda_op.cu:1065-1269),
 *                                         here an exhaustive search with the same closest-point-on-triangle
 *                                         evaluation (bvh_cuda_op.cu:209-281); ties go to the lowest face index;
 *   get_id                             -- include/utils.h:468-481 (first of the <= 2 adjacent tets that passes)
 *   point_in_tet / same_side           -- include/utils.h:394-423
 *   barycentric                        -- include/utils.h:425-451 (`1.0 / triple` is a double divide rounded to float)
 *   miss -> barys 0, tetra_id -1, active -1   -- src/sampler_points_kernel.cu:299-305
 * Pinned against the reference's own header through oracle/_ref/libtetmath_ref.so (tests/test_oracle_tet.py) and the
 * committed vectors tests/golden/tet_golden.npz generated from it.
 */
#include <math.h>
#include <stdint.h>

typedef struct { float x, y, z; } v3;
static inline v3 sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline v3 add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline v3 mul(float s, v3 a) { v3 r = {s * a.x, s * a.y, s * a.z}; return r; }
static inline float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 cross(v3 a, v3 b) { v3 r = {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; return r; }
static inline v3 ld(const float *p) { v3 r = {p[0], p[1], p[2]}; return r; }

static float point_triangle_dist2(v3 p, v3 a, v3 b, v3 c) {
    v3 ab = sub(b, a), ac = sub(c, a), ap = sub(p, a);
    float d1 = dot(ab, ap), d2 = dot(ac, ap);
    if (d1 <= 0.f && d2 <= 0.f) return dot(ap, ap);
    v3 bp = sub(p, b);
    float d3 = dot(ab, bp), d4 = dot(ac, bp);
    if (d3 >= 0.f && d4 <= d3) return dot(bp, bp);
    float vc = d1 * d4 - d3 * d2;
    if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { float v = d1 / (d1 - d3); v3 d = sub(p, add(a, mul(v, ab))); return dot(d, d); }
    v3 cp = sub(p, c);
    float d5 = dot(ab, cp), d6 = dot(ac, cp);
    if (d6 >= 0.f && d5 <= d6) return dot(cp, cp);
    float vb = d5 * d2 - d1 * d6;
    if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { float w = d2 / (d2 - d6); v3 d = sub(p, add(a, mul(w, ac))); return dot(d, d); }
    float va = d3 * d6 - d5 * d4;
    if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) {
        float w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
        v3 d = sub(p, add(b, mul(w, sub(c, b))));
        return dot(d, d);
    }
    float denom = 1.f / (va + vb + vc), v = vb * denom, w = vc * denom;
    v3 d = sub(p, add(add(a, mul(v, ab)), mul(w, ac)));
    return dot(d, d);
}

static int same_side(v3 v1, v3 v2, v3 v3_, v3 v4, v3 p) {
    v3 normal = cross(sub(v2, v1), sub(v3_, v1));
    return signbit(dot(normal, sub(v4, v1))) == signbit(dot(normal, sub(p, v1)));
}
int orc_point_in_tet(const float *t, const float *p_) {
    v3 v1 = ld(t), v2 = ld(t + 3), v3_ = ld(t + 6), v4 = ld(t + 9), p = ld(p_);
    return same_side(v1, v2, v3_, v4, p) && same_side(v2, v3_, v4, v1, p) && same_side(v3_, v4, v1, v2, p) &&
           same_side(v4, v1, v2, v3_, p);
}
void orc_barycentric(const float *t, const float *p_, float *out) {
    v3 a = ld(t), b = ld(t + 3), c = ld(t + 6), d = ld(t + 9), p = ld(p_);
    v3 vap = sub(p, a), vbp = sub(p, b), vab = sub(b, a), vac = sub(c, a), vad = sub(d, a), vbc = sub(c, b), vbd = sub(d, b);
    float va6 = dot(vbp, cross(vbd, vbc)), vb6 = dot(vap, cross(vac, vad));
    float vc6 = dot(vap, cross(vad, vab)), vd6 = dot(vap, cross(vab, vac));
    float v6 = (float)(1.0 / (double)dot(vab, cross(vac, vad)));
    out[0] = va6 * v6; out[1] = vb6 * v6; out[2] = vc6 * v6; out[3] = vd6 * v6;
}

void orc_compute_bary(int N, int T, int F, const float *points, const float *tetras, const float *triangles,
                      const int32_t *tri_to_tetra, float *barys, int32_t *tetra_id, int32_t *active,
                      int32_t *closest_face) {
    (void)T;
#pragma omp parallel for schedule(static)
    for (int n = 0; n < N; n++) {
        v3 p = ld(points + 3 * n);
        float best = INFINITY; int bf = -1;
        for (int f = 0; f < F; f++) {
            const float *t = triangles + 9 * (long)f;
            float d = point_triangle_dist2(p, ld(t), ld(t + 3), ld(t + 6));
            if (d < best) { best = d; bf = f; }
        }
        if (closest_face) closest_face[n] = bf;
        int id = -1;
        if (bf >= 0) {
            int a = tri_to_tetra[2 * bf], b = tri_to_tetra[2 * bf + 1];
            if (a != -1 && orc_point_in_tet(tetras + 12 * (long)a, points + 3 * n)) id = a;
            else if (b != -1 && orc_point_in_tet(tetras + 12 * (long)b, points + 3 * n)) id = b;
        }
        if (id == -1) { barys[4 * n] = barys[4 * n + 1] = barys[4 * n + 2] = barys[4 * n + 3] = 0.f; tetra_id[n] = -1; active[n] = -1; continue; }
        orc_barycentric(tetras + 12 * (long)id, points + 3 * n, barys + 4 * n);
        tetra_id[n] = id; active[n] = 1;
    }
}
