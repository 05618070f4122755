// oracle/ref_tetmath_shim.cu -- TEST INFRASTRUCTURE ONLY.
// Thin extern "C" window onto the REFERENCE's own host-callable tet math, compiled from the header where it lies:
//   /root/reference/submodules/tetra-sampler/include/utils.h:394-451  (same_side, point_in_tet, barycentric)
// Nothing of the reference is copied into this repository; `make ref` builds oracle/_ref/libtetmath_ref.so, which is
// git-ignored but travels to the GPU box.  Used to pin oracle/tet_oracle.c and to generate tests/golden/tet_golden.npz.
#include <cuda_runtime.h>
#include <cstdint>
#include <utility>
#include "utils.h"

extern "C" void ref_barycentric(int n, const float *points, const float *tets /* [n,4,3] */, float *out /* [n,4] */) {
    for (int i = 0; i < n; i++) {
        const float *t = tets + 12 * i;
        Tetra tet;
        tet.v0 = make_float3(t[0], t[1], t[2]); tet.v1 = make_float3(t[3], t[4], t[5]);
        tet.v2 = make_float3(t[6], t[7], t[8]); tet.v3 = make_float3(t[9], t[10], t[11]);
        float4 b = barycentric(make_float3(points[3 * i], points[3 * i + 1], points[3 * i + 2]), tet);
        out[4 * i] = b.x; out[4 * i + 1] = b.y; out[4 * i + 2] = b.z; out[4 * i + 3] = b.w;
    }
}

extern "C" void ref_point_in_tet(int n, const float *points, const float *tets, int32_t *out) {
    for (int i = 0; i < n; i++) {
        const float *t = tets + 12 * i;
        Tetra tet;
        tet.v0 = make_float3(t[0], t[1], t[2]); tet.v1 = make_float3(t[3], t[4], t[5]);
        tet.v2 = make_float3(t[6], t[7], t[8]); tet.v3 = make_float3(t[9], t[10], t[11]);
        out[i] = point_in_tet(tet, make_float3(points[3 * i], points[3 * i + 1], points[3 * i + 2])) ? 1 : 0;
    }
}
