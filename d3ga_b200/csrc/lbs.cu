// lbs.cu -- SURVEY.md section 8(f) rank 4: cage-vertex linear blend skinning, the step that produces `tetpoints`
// immediately before the deformation kernel.
//
// Replaces, per frame and per garment:
//   * ParameterTransform.forward + solve_skeleton_state + states_to_matrix (lbsmodel/body_model.py:39-47, :311-387) --
//     in the reference a Python loop over the joints with ~10 torch launches per joint (over a thousand launches for a
//     ~100-joint skeleton) -- by ONE kernel: one CTA per pose, joints solved level by level in shared memory;
//   * LinearBlendSkinning.skinning (body_model.py:208-234) + LBSModule.pose's global scaling (:669-675) + Blueman.get's
//     root-relative transform, metre conversion and centre-of-mass shift (lib/blueman.py:101-116,139-158) by ONE kernel,
//     one thread per cage vertex, with its backward (gradient to the template vertices / `delta`);
//   * LinearBlendSkinning.unskinning (body_model.py:248-283; init only) and Smplman.deform (lib/smplman.py:155-171).
// Everything is latency-bound (V <= ~20 k vertices, ~100 joints): the win is launch count, not bandwidth.
// Quaternions are (x, y, z, w) like lbsmodel/quaternion.py.  Plain fp32; sincosf / exp2f are the accurate versions.
#include "common.cuh"

namespace d3ga {
namespace {

constexpr int LBS_MAX_JOINTS = 1024;
constexpr int LBS_THREADS = 256;

struct Quat { float x, y, z, w; };
struct Vec3 { float x, y, z; };

__device__ __forceinline__ Quat qmul(const Quat &q, const Quat &r) {      // quaternion.py:169-236
    Quat o;
    o.x = q.x * r.w + q.y * r.z - q.z * r.y + q.w * r.x;
    o.y = -q.x * r.z + q.y * r.w + q.z * r.x + q.w * r.y;
    o.z = q.x * r.y - q.y * r.x + q.z * r.w + q.w * r.z;
    o.w = -q.x * r.x - q.y * r.y - q.z * r.z + q.w * r.w;
    return o;
}
__device__ __forceinline__ Vec3 cross3(const Vec3 &a, const Vec3 &b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ Vec3 qrot(const Quat &q, const Vec3 &v) {      // quaternion.py:238-252
    const Vec3 a = {q.x, q.y, q.z};
    const Vec3 av = cross3(a, v), aav = cross3(a, av);
    return {v.x + 2.f * (av.x * q.w + aav.x), v.y + 2.f * (av.y * q.w + aav.y), v.z + 2.f * (av.z * q.w + aav.z)};
}
__device__ __forceinline__ Quat qinv(const Quat &q) {                     // quaternion.py:254-265
    const float r = 1.f / (q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
    return {-q.x * r, -q.y * r, -q.z * r, q.w * r};
}
__device__ __forceinline__ Quat qfrom_xyz(float rx, float ry, float rz) {  // quaternion.py:267-304
    float s0, c0, s1, c1, s2, c2;
    sincosf(rx * -0.5f, &s0, &c0);
    sincosf(ry * 0.5f, &s1, &c1);
    sincosf(rz * 0.5f, &s2, &c2);
    return {-s0 * (c1 * c2) - c0 * (s1 * s2), c0 * (s1 * c2) - s0 * (c1 * s2), c0 * (c1 * s2) + s0 * (s1 * c2),
            c0 * (c1 * c2) - s0 * (s1 * s2)};
}

// states_to_matrix (body_model.py:350-387) for one joint: target o bind^-1 as a row-major [3][4] matrix.
__device__ __forceinline__ void state_to_matrix(const float *bd, const float *g, float *m) {
    const Quat br = qinv(Quat{bd[3], bd[4], bd[5], bd[6]});
    const float bs = 1.f / bd[7];
    Vec3 bt = qrot(br, Vec3{-bd[0], -bd[1], -bd[2]});
    bt.x *= bs; bt.y *= bs; bt.z *= bs;
    const Quat tq = {g[3], g[4], g[5], g[6]};
    const Quat tr = qmul(tq, br);
    const float ts = g[7] * bs;
    Vec3 tt = qrot(tq, Vec3{bt.x * g[7], bt.y * g[7], bt.z * g[7]});
    tt.x += g[0]; tt.y += g[1]; tt.z += g[2];
    const float twx = 2.f * tr.x * tr.w, twy = 2.f * tr.y * tr.w, twz = 2.f * tr.z * tr.w;
    const float txx = 2.f * tr.x * tr.x, txy = 2.f * tr.y * tr.x, txz = 2.f * tr.z * tr.x;
    const float tyy = 2.f * tr.y * tr.y, tyz = 2.f * tr.z * tr.y, tzz = 2.f * tr.z * tr.z;
    m[0] = (1.f - (tyy + tzz)) * ts; m[1] = (txy - twz) * ts;         m[2] = (txz + twy) * ts;          m[3] = tt.x;
    m[4] = (txy + twz) * ts;         m[5] = (1.f - (txx + tzz)) * ts; m[6] = (tyz - twx) * ts;          m[7] = tt.y;
    m[8] = (txz - twy) * ts;         m[9] = (tyz + twx) * ts;         m[10] = (1.f - (txx + tyy)) * ts; m[11] = tt.z;
}

__global__ void __launch_bounds__(LBS_THREADS) lbs_states_to_matrix_kernel(int n, int NJ, const float *__restrict__ bind_state,
                                                                           const float *__restrict__ states,
                                                                           float *__restrict__ mats) {
    const int i = blockIdx.x * LBS_THREADS + threadIdx.x;        // (pose, joint) pair
    if (i >= n) return;
    float g[8], m[12];
#pragma unroll
    for (int k = 0; k < 8; k++) g[k] = states[(size_t)i * 8 + k];
    state_to_matrix(bind_state + 8 * (i % NJ), g, m);
#pragma unroll
    for (int k = 0; k < 12; k++) mats[(size_t)i * 12 + k] = m[k];
}

// One CTA per pose.  Shared: joint parameters [7 NJ] then reused; states [8 NJ]; level [NJ].
__global__ void __launch_bounds__(LBS_THREADS)
lbs_skeleton_kernel(int NJ, int batch, int n_pose, int n_scale, const float *__restrict__ pose,
                    const float *__restrict__ scales, int64_t scales_stride, const int32_t *__restrict__ row_ptr,
                    const int32_t *__restrict__ col_idx, const float *__restrict__ vals, const float *__restrict__ offsets,
                    const float *__restrict__ joint_offset, const float *__restrict__ joint_rotation,
                    const int32_t *__restrict__ joint_parents, const float *__restrict__ bind_state,
                    float *__restrict__ states_out, float *__restrict__ mats_out) {
    extern __shared__ float smem[];
    float *jp = smem;                          // [7 NJ] transformed parameters, later local (t, q, s) in place + 1
    float *st = smem + 8 * NJ;                 // [8 NJ] global states
    volatile int *level = reinterpret_cast<volatile int *>(smem + 16 * NJ);
    const int b = blockIdx.x;                  // CTAs [batch, 2 batch) repeat the poses with all scale parameters zero
    const float *p = pose + (size_t)(b % batch) * n_pose;
    const float *sc = (b < batch && scales) ? scales + (size_t)b * scales_stride : nullptr;

    // ParameterTransform.forward: sparse rows of `transform` (the host builds CSR from the dense matrix once)
    for (int r = threadIdx.x; r < 7 * NJ; r += LBS_THREADS) {
        float acc = 0.f;
        for (int e = row_ptr[r]; e < row_ptr[r + 1]; e++) {
            const int c = col_idx[e];
            const float x = c < n_pose ? __ldg(p + c) : (sc && c - n_pose < n_scale ? __ldg(sc + c - n_pose) : 0.f);
            acc += vals[e] * x;
        }
        jp[(r / 7) * 8 + (r % 7)] = acc + offsets[r];
    }
    for (int j = threadIdx.x; j < NJ; j += LBS_THREADS) level[j] = 0x7fffffff;
    __syncthreads();

    // local transforms (body_model.py:318-325): lt = t + offset, lr = prerotation * fromXYZ(r), ls = 2^s
    for (int j = threadIdx.x; j < NJ; j += LBS_THREADS) {
        float *q = jp + 8 * j;
        const Quat pre = {joint_rotation[4 * j], joint_rotation[4 * j + 1], joint_rotation[4 * j + 2], joint_rotation[4 * j + 3]};
        const Quat lr = qmul(pre, qfrom_xyz(q[3], q[4], q[5]));
        const float ls = exp2f(q[6]);
        q[0] += joint_offset[3 * j];
        q[1] += joint_offset[3 * j + 1];
        q[2] += joint_offset[3 * j + 2];
        q[3] = lr.x; q[4] = lr.y; q[5] = lr.z; q[6] = lr.w; q[7] = ls;
    }
    __syncthreads();

    // forward kinematics (body_model.py:327-347), one tree level per iteration: a joint is solved in the first
    // iteration that starts after its parent was solved.
    for (int it = 0; it <= NJ; it++) {
        bool pending = false;
        for (int j = threadIdx.x; j < NJ; j += LBS_THREADS) {
            if (level[j] <= it) continue;
            const int par = joint_parents[j];
            const float *l = jp + 8 * j;
            float *g = st + 8 * j;
            if (par < 0 || par >= NJ) {
#pragma unroll
                for (int k = 0; k < 8; k++) g[k] = l[k];
                level[j] = it;
            } else if (level[par] < it) {
                const float *ps = st + 8 * par;
                const Quat pq = {ps[3], ps[4], ps[5], ps[6]};
                const Quat gr = qmul(pq, Quat{l[3], l[4], l[5], l[6]});
                const Vec3 gt = qrot(pq, Vec3{l[0] * ps[7], l[1] * ps[7], l[2] * ps[7]});
                g[0] = gt.x + ps[0]; g[1] = gt.y + ps[1]; g[2] = gt.z + ps[2];
                g[3] = gr.x; g[4] = gr.y; g[5] = gr.z; g[6] = gr.w;
                g[7] = ps[7] * l[7];
                level[j] = it;
            } else {
                pending = true;
            }
        }
        if (!__syncthreads_or(pending)) break;
    }

    for (int j = threadIdx.x; j < NJ; j += LBS_THREADS) {
        const float *g = st + 8 * j;
        if (states_out) {
#pragma unroll
            for (int k = 0; k < 8; k++) states_out[((size_t)b * NJ + j) * 8 + k] = g[k];
        }
        if (!mats_out) continue;
        state_to_matrix(bind_state + 8 * j, g, mats_out + ((size_t)b * NJ + j) * 12);
    }
}

// ---- skinning ---------------------------------------------------------------------------------------------------------
struct SkinArgs {
    int V, K, NJ;
    const float *verts;  int64_t verts_stride;        // floats between poses (0 = one template for all)
    const float *delta;  int64_t delta_stride;  float delta_unit;   // v = (t / unit + delta) * unit   (blueman.py:105-106)
    const int64_t *skin_indices;
    const float *skin_weights;
    const float *mats;                                   // [B, NJ, 12]
    const float *global_scaling;                         // [3] or null
    const float *root_mats;  int root_joint;             // [B, NJ, 12] of the zero-scale solve, or null = no root removal
    const float *pre_mul;                                // [16] row-major 4x4 right factor (rot180), or null
    float unit_div;                                      // geom / unit_div (1000: mm -> m)
    const float *shift;                                  // [3] centre of mass, or null
};

// post transform of Blueman.get: x -> Rinv (x / unit_div) + tinv + shift, with [Rinv | tinv] = inverse([R_root | t_root /
// unit_div] @ pre_mul).  Computed redundantly by one thread per CTA (a 3x3 inverse) and broadcast through shared memory.
__device__ void root_post_transform(const SkinArgs &a, int b, float (&post)[12], float *rt_out) {
    float A[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}}, t[3] = {0, 0, 0};
    if (a.root_mats) {
        const float *m = a.root_mats + ((size_t)b * a.NJ + a.root_joint) * 12;
        float R[3][3], tr[3];
#pragma unroll
        for (int r = 0; r < 3; r++) {
#pragma unroll
            for (int c = 0; c < 3; c++) R[r][c] = m[4 * r + c];
            tr[r] = m[4 * r + 3] / a.unit_div;
        }
        if (a.pre_mul) {                                  // [R | t] @ P for a 4x4 P with last row (0,0,0,1)
            const float *P = a.pre_mul;
#pragma unroll
            for (int r = 0; r < 3; r++) {
#pragma unroll
                for (int c = 0; c < 3; c++) A[r][c] = R[r][0] * P[c] + R[r][1] * P[4 + c] + R[r][2] * P[8 + c];
                t[r] = R[r][0] * P[3] + R[r][1] * P[7] + R[r][2] * P[11] + tr[r];
            }
        } else {
#pragma unroll
            for (int r = 0; r < 3; r++) {
#pragma unroll
                for (int c = 0; c < 3; c++) A[r][c] = R[r][c];
                t[r] = tr[r];
            }
        }
    }
    // inverse of the affine map [A | t]
    const float c00 = A[1][1] * A[2][2] - A[1][2] * A[2][1], c01 = A[1][2] * A[2][0] - A[1][0] * A[2][2],
                c02 = A[1][0] * A[2][1] - A[1][1] * A[2][0];
    const float det = A[0][0] * c00 + A[0][1] * c01 + A[0][2] * c02;
    const float id = 1.f / det;
    float I[3][3];
    I[0][0] = c00 * id; I[1][0] = c01 * id; I[2][0] = c02 * id;
    I[0][1] = (A[0][2] * A[2][1] - A[0][1] * A[2][2]) * id;
    I[1][1] = (A[0][0] * A[2][2] - A[0][2] * A[2][0]) * id;
    I[2][1] = (A[0][1] * A[2][0] - A[0][0] * A[2][1]) * id;
    I[0][2] = (A[0][1] * A[1][2] - A[0][2] * A[1][1]) * id;
    I[1][2] = (A[0][2] * A[1][0] - A[0][0] * A[1][2]) * id;
    I[2][2] = (A[0][0] * A[1][1] - A[0][1] * A[1][0]) * id;
#pragma unroll
    for (int r = 0; r < 3; r++) {
#pragma unroll
        for (int c = 0; c < 3; c++) post[4 * r + c] = I[r][c];
        post[4 * r + 3] = -(I[r][0] * t[0] + I[r][1] * t[1] + I[r][2] * t[2]);
    }
    if (rt_out) {
#pragma unroll
        for (int k = 0; k < 12; k++) rt_out[(size_t)b * 16 + k] = post[k];
        rt_out[(size_t)b * 16 + 12] = 0.f; rt_out[(size_t)b * 16 + 13] = 0.f; rt_out[(size_t)b * 16 + 14] = 0.f;
        rt_out[(size_t)b * 16 + 15] = 1.f;
    }
}

__device__ __forceinline__ void load_vertex(const SkinArgs &a, int b, int v, float (&x)[3]) {
    const float *t = a.verts + (size_t)b * a.verts_stride + 3 * (size_t)v;
    if (a.delta) {
        const float *d = a.delta + (size_t)b * a.delta_stride + 3 * (size_t)v;
#pragma unroll
        for (int c = 0; c < 3; c++) x[c] = (t[c] / a.delta_unit + d[c]) * a.delta_unit;
    } else {
#pragma unroll
        for (int c = 0; c < 3; c++) x[c] = t[c];
    }
}

__global__ void __launch_bounds__(LBS_THREADS) lbs_skin_forward_kernel(SkinArgs a, float *__restrict__ out,
                                                                       float *__restrict__ rt_out) {
    __shared__ float post[12];
    const int b = blockIdx.y;
    if (threadIdx.x == 0) {
        float p[12];
        root_post_transform(a, b, p, blockIdx.x == 0 ? rt_out : nullptr);
#pragma unroll
        for (int k = 0; k < 12; k++) post[k] = p[k];
    }
    __syncthreads();
    const int v = blockIdx.x * LBS_THREADS + threadIdx.x;
    if (v >= a.V) return;
    float x[3];
    load_vertex(a, b, v, x);
    float acc[3] = {0.f, 0.f, 0.f};
    const float4 *mats = reinterpret_cast<const float4 *>(a.mats) + (size_t)b * a.NJ * 3;
    for (int k = 0; k < a.K; k++) {
        const float w = a.skin_weights[(size_t)v * a.K + k];
        const int64_t j = a.skin_indices[(size_t)v * a.K + k];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float4 m = __ldg(mats + 3 * j + r);
            acc[r] += (m.x * x[0] + m.y * x[1] + m.z * x[2] + m.w) * w;
        }
    }
#pragma unroll
    for (int r = 0; r < 3; r++) acc[r] = (a.global_scaling ? acc[r] * a.global_scaling[r] : acc[r]) / a.unit_div;
#pragma unroll
    for (int r = 0; r < 3; r++)
        out[((size_t)b * a.V + v) * 3 + r] = post[4 * r] * acc[0] + post[4 * r + 1] * acc[1] + post[4 * r + 2] * acc[2] +
                                             post[4 * r + 3] + (a.shift ? a.shift[r] : 0.f);
}

// gradient to the (effective) template vertex: dv = sum_k w_k M_k[:, :3]^T (gs / unit_div) Rinv^T dout, times grad_scale
// (delta_unit when the caller differentiates with respect to `delta`).
__global__ void __launch_bounds__(LBS_THREADS) lbs_skin_backward_kernel(SkinArgs a, const float *__restrict__ dL_dout,
                                                                        float grad_scale, float *__restrict__ dL_dverts) {
    __shared__ float post[12];
    const int b = blockIdx.y;
    if (threadIdx.x == 0) {
        float p[12];
        root_post_transform(a, b, p, nullptr);
#pragma unroll
        for (int k = 0; k < 12; k++) post[k] = p[k];
    }
    __syncthreads();
    const int v = blockIdx.x * LBS_THREADS + threadIdx.x;
    if (v >= a.V) return;
    const float *g = dL_dout + ((size_t)b * a.V + v) * 3;
    float d[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        d[c] = (post[c] * g[0] + post[4 + c] * g[1] + post[8 + c] * g[2]) / a.unit_div;
        if (a.global_scaling) d[c] *= a.global_scaling[c];
    }
    float dv[3] = {0.f, 0.f, 0.f};
    const float4 *mats = reinterpret_cast<const float4 *>(a.mats) + (size_t)b * a.NJ * 3;
    for (int k = 0; k < a.K; k++) {
        const float w = a.skin_weights[(size_t)v * a.K + k];
        const int64_t j = a.skin_indices[(size_t)v * a.K + k];
        const float4 m0 = __ldg(mats + 3 * j), m1 = __ldg(mats + 3 * j + 1), m2 = __ldg(mats + 3 * j + 2);
        dv[0] += w * (m0.x * d[0] + m1.x * d[1] + m2.x * d[2]);
        dv[1] += w * (m0.y * d[0] + m1.y * d[1] + m2.y * d[2]);
        dv[2] += w * (m0.z * d[0] + m1.z * d[1] + m2.z * d[2]);
    }
#pragma unroll
    for (int c = 0; c < 3; c++) dL_dverts[((size_t)b * a.V + v) * 3 + c] = dv[c] * grad_scale;
}

// LinearBlendSkinning.unskinning: blend the 3x4 matrices, invert the affine map, apply it.
__global__ void __launch_bounds__(LBS_THREADS)
lbs_unskin_kernel(int V, int K, int NJ, const float *__restrict__ verts, const int64_t *__restrict__ skin_indices,
                  const float *__restrict__ skin_weights, const float *__restrict__ mats, float *__restrict__ out) {
    const int b = blockIdx.y;
    const int v = blockIdx.x * LBS_THREADS + threadIdx.x;
    if (v >= V) return;
    float M[12];
#pragma unroll
    for (int k = 0; k < 12; k++) M[k] = 0.f;
    const float4 *mp = reinterpret_cast<const float4 *>(mats) + (size_t)b * NJ * 3;
    for (int k = 0; k < K; k++) {
        const float w = skin_weights[(size_t)v * K + k];
        const int64_t j = skin_indices[(size_t)v * K + k];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float4 m = __ldg(mp + 3 * j + r);
            M[4 * r] += m.x * w; M[4 * r + 1] += m.y * w; M[4 * r + 2] += m.z * w; M[4 * r + 3] += m.w * w;
        }
    }
    const float *x = verts + ((size_t)b * V + v) * 3;
    const float y[3] = {x[0] - M[3], x[1] - M[7], x[2] - M[11]};
    const float c00 = M[5] * M[10] - M[6] * M[9], c01 = M[6] * M[8] - M[4] * M[10], c02 = M[4] * M[9] - M[5] * M[8];
    const float id = 1.f / (M[0] * c00 + M[1] * c01 + M[2] * c02);
    float *o = out + ((size_t)b * V + v) * 3;
    o[0] = (c00 * y[0] + (M[2] * M[9] - M[1] * M[10]) * y[1] + (M[1] * M[6] - M[2] * M[5]) * y[2]) * id;
    o[1] = (c01 * y[0] + (M[0] * M[10] - M[2] * M[8]) * y[1] + (M[2] * M[4] - M[0] * M[6]) * y[2]) * id;
    o[2] = (c02 * y[0] + (M[1] * M[8] - M[0] * M[9]) * y[1] + (M[0] * M[5] - M[1] * M[4]) * y[2]) * id;
}

// ---- Smplman.deform: dense weights over J joints, 4x4 joint matrices, Rodrigues global rotation -------------------------
__device__ void rodrigues(const float *rh, float (&R)[9]) {               // tetra_sampler/lbs.py:272-302
    const float ax = rh[0] + 1e-8f, ay = rh[1] + 1e-8f, az = rh[2] + 1e-8f;
    const float angle = sqrtf(ax * ax + ay * ay + az * az);
    const float x = rh[0] / angle, y = rh[1] / angle, z = rh[2] / angle;
    float s, c;
    sincosf(angle, &s, &c);
    const float K[9] = {0.f, -z, y, z, 0.f, -x, -y, x, 0.f};
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const float kk = K[3 * r] * K[q] + K[3 * r + 1] * K[3 + q] + K[3 * r + 2] * K[6 + q];
            R[3 * r + q] = (r == q ? 1.f : 0.f) + s * K[3 * r + q] + (1.f - c) * kk;
        }
}

template <bool BACKWARD>
__global__ void __launch_bounds__(LBS_THREADS)
lbs_skin_dense_kernel(int V, int Vd, int J, const float *__restrict__ verts, const float *__restrict__ delta,
                      const float *__restrict__ weights, const int64_t *__restrict__ nn_ids, const float *__restrict__ A,
                      const float *__restrict__ Bs, const float *__restrict__ Rh, const float *__restrict__ Th,
                      const float *__restrict__ dL_dout, float *__restrict__ out) {
    extern __shared__ float sA[];                                         // [J, 12] top three rows of A
    __shared__ float R[9];
    const int b = blockIdx.y;
    for (int i = threadIdx.x; i < J * 12; i += LBS_THREADS) sA[i] = A[((size_t)b * J + i / 12) * 16 + i % 12];
    if (threadIdx.x == 0) {
        float r[9];
        rodrigues(Rh + 3 * b, r);
#pragma unroll
        for (int k = 0; k < 9; k++) R[k] = r[k];
    }
    __syncthreads();
    const int v = blockIdx.x * LBS_THREADS + threadIdx.x;
    if (v >= V) return;
    const int64_t nn = nn_ids[v];
    float T[12];
#pragma unroll
    for (int k = 0; k < 12; k++) T[k] = 0.f;
    const float *w = weights + (size_t)nn * J;
    for (int j = 0; j < J; j++) {
        const float wj = __ldg(w + j);
        if (wj == 0.f) continue;
#pragma unroll
        for (int k = 0; k < 12; k++) T[k] += wj * sA[12 * j + k];
    }
    if (!BACKWARD) {
        float x[3];
#pragma unroll
        for (int c = 0; c < 3; c++)
            x[c] = (delta ? verts[3 * (size_t)v + c] + delta[3 * (size_t)v + c] : verts[3 * (size_t)v + c]) +
                   Bs[((size_t)b * Vd + nn) * 3 + c];
        float g[3];
#pragma unroll
        for (int r = 0; r < 3; r++) g[r] = T[4 * r] * x[0] + T[4 * r + 1] * x[1] + T[4 * r + 2] * x[2] + T[4 * r + 3];
#pragma unroll
        for (int r = 0; r < 3; r++)
            out[((size_t)b * V + v) * 3 + r] = g[0] * R[3 * r] + g[1] * R[3 * r + 1] + g[2] * R[3 * r + 2] + Th[3 * b + r];
    } else {
        const float *go = dL_dout + ((size_t)b * V + v) * 3;
        float dg[3];
#pragma unroll
        for (int c = 0; c < 3; c++) dg[c] = R[c] * go[0] + R[3 + c] * go[1] + R[6 + c] * go[2];
#pragma unroll
        for (int c = 0; c < 3; c++) out[((size_t)b * V + v) * 3 + c] = T[c] * dg[0] + T[4 + c] * dg[1] + T[8 + c] * dg[2];
    }
}

}  // namespace
}  // namespace d3ga

using namespace d3ga;

extern "C" int d3ga_lbs_skeleton_forward(int batch, int n_joints, int n_pose, int n_scale, const float *pose,
                                         const float *scales, int64_t scales_stride, int with_zero_scale_copy,
                                         const int32_t *row_ptr, const int32_t *col_idx, const float *vals,
                                         const float *offsets, const float *joint_offset, const float *joint_rotation,
                                         const int32_t *joint_parents, const float *bind_state, float *states, float *mats,
                                         d3ga_stream_t stream) {
    if (batch < 0 || n_joints <= 0 || n_pose < 0 || n_scale < 0 || n_pose + n_scale <= 0) return D3GA_EINVAL;
    if (n_joints > LBS_MAX_JOINTS) return D3GA_ELIMIT;
    if (batch == 0) return D3GA_OK;
    // col_idx / vals may be null for a transform without a single non-zero (row_ptr all zero)
    if ((n_pose && !pose) || !row_ptr || (!col_idx != !vals) || !offsets || !joint_offset || !joint_rotation || !joint_parents)
        return D3GA_EINVAL;
    if ((!states && !mats) || (mats && !bind_state)) return D3GA_EINVAL;
    ProfScope prof(K_LBS_SKELETON, (cudaStream_t)stream);
    const size_t smem = (size_t)n_joints * 17 * sizeof(float);
    if (smem > 48 * 1024)
        D3GA_CUDA(cudaFuncSetAttribute(lbs_skeleton_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lbs_skeleton_kernel<<<with_zero_scale_copy ? 2 * batch : batch, LBS_THREADS, smem, (cudaStream_t)stream>>>(
        n_joints, batch, n_pose, n_scale, pose, scales, scales_stride, row_ptr, col_idx, vals, offsets, joint_offset,
        joint_rotation, joint_parents, bind_state, states, mats);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_lbs_states_to_matrix(int batch, int n_joints, const float *bind_state, const float *states, float *mats,
                                         d3ga_stream_t stream) {
    if (batch < 0 || n_joints <= 0) return D3GA_EINVAL;
    if (batch == 0) return D3GA_OK;
    if (!bind_state || !states || !mats) return D3GA_EINVAL;
    ProfScope prof(K_LBS_SKELETON, (cudaStream_t)stream);
    const int n = batch * n_joints;
    lbs_states_to_matrix_kernel<<<(n + LBS_THREADS - 1) / LBS_THREADS, LBS_THREADS, 0, (cudaStream_t)stream>>>(
        n, n_joints, bind_state, states, mats);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

static int check_skin(int batch, int V, int K, int NJ, const SkinArgs &a) {
    if (batch < 0 || V < 0 || K <= 0 || NJ <= 0) return D3GA_EINVAL;
    if (batch == 0 || V == 0) return 1;
    if (!a.verts || !a.skin_indices || !a.skin_weights || !a.mats) return D3GA_EINVAL;
    if (a.root_mats && (a.root_joint < 0 || a.root_joint >= NJ)) return D3GA_EINVAL;
    if (!(a.unit_div != 0.f) || (a.delta && !(a.delta_unit != 0.f))) return D3GA_EINVAL;
    return D3GA_OK;
}

static SkinArgs make_skin_args(int V, int K, int NJ, const float *verts, int64_t verts_stride, const float *delta,
                               int64_t delta_stride, float delta_unit, const int64_t *skin_indices,
                               const float *skin_weights, const float *mats, const float *global_scaling,
                               const float *root_mats, int root_joint, const float *pre_mul, float unit_div,
                               const float *shift) {
    SkinArgs a;
    a.V = V; a.K = K; a.NJ = NJ;
    a.verts = verts; a.verts_stride = verts_stride;
    a.delta = delta; a.delta_stride = delta_stride; a.delta_unit = delta_unit;
    a.skin_indices = skin_indices; a.skin_weights = skin_weights; a.mats = mats; a.global_scaling = global_scaling;
    a.root_mats = root_mats; a.root_joint = root_joint; a.pre_mul = pre_mul; a.unit_div = unit_div; a.shift = shift;
    return a;
}

extern "C" int d3ga_lbs_skin_forward(int batch, int V, int K, int n_joints, const float *verts, int64_t verts_stride,
                                     const float *delta, int64_t delta_stride, float delta_unit,
                                     const int64_t *skin_indices, const float *skin_weights, const float *mats,
                                     const float *global_scaling, const float *root_mats, int root_joint,
                                     const float *pre_mul, float unit_div, const float *shift, float *out, float *rt_out,
                                     d3ga_stream_t stream) {
    const SkinArgs a = make_skin_args(V, K, n_joints, verts, verts_stride, delta, delta_stride, delta_unit, skin_indices,
                                      skin_weights, mats, global_scaling, root_mats, root_joint, pre_mul, unit_div, shift);
    const int rc = check_skin(batch, V, K, n_joints, a);
    if (rc) return rc > 0 ? D3GA_OK : rc;
    if (!out) return D3GA_EINVAL;
    ProfScope prof(K_LBS_SKIN_FWD, (cudaStream_t)stream);
    const dim3 grid((V + LBS_THREADS - 1) / LBS_THREADS, batch);
    lbs_skin_forward_kernel<<<grid, LBS_THREADS, 0, (cudaStream_t)stream>>>(a, out, rt_out);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_lbs_skin_backward(int batch, int V, int K, int n_joints, const int64_t *skin_indices,
                                      const float *skin_weights, const float *mats, const float *global_scaling,
                                      const float *root_mats, int root_joint, const float *pre_mul, float unit_div,
                                      const float *dL_dout, float grad_scale, float *dL_dverts, d3ga_stream_t stream) {
    // the backward needs no vertex positions (the map is affine in them); a non-null dummy satisfies the shared check
    const SkinArgs a = make_skin_args(V, K, n_joints, mats, 0, nullptr, 0, 1.f, skin_indices, skin_weights, mats,
                                      global_scaling, root_mats, root_joint, pre_mul, unit_div, nullptr);
    const int rc = check_skin(batch, V, K, n_joints, a);
    if (rc) return rc > 0 ? D3GA_OK : rc;
    if (!dL_dout || !dL_dverts) return D3GA_EINVAL;
    ProfScope prof(K_LBS_SKIN_BWD, (cudaStream_t)stream);
    const dim3 grid((V + LBS_THREADS - 1) / LBS_THREADS, batch);
    lbs_skin_backward_kernel<<<grid, LBS_THREADS, 0, (cudaStream_t)stream>>>(a, dL_dout, grad_scale, dL_dverts);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_lbs_unskin_forward(int batch, int V, int K, int n_joints, const float *verts,
                                       const int64_t *skin_indices, const float *skin_weights, const float *mats,
                                       float *out, d3ga_stream_t stream) {
    if (batch < 0 || V < 0 || K <= 0 || n_joints <= 0) return D3GA_EINVAL;
    if (batch == 0 || V == 0) return D3GA_OK;
    if (!verts || !skin_indices || !skin_weights || !mats || !out) return D3GA_EINVAL;
    ProfScope prof(K_LBS_UNSKIN, (cudaStream_t)stream);
    const dim3 grid((V + LBS_THREADS - 1) / LBS_THREADS, batch);
    lbs_unskin_kernel<<<grid, LBS_THREADS, 0, (cudaStream_t)stream>>>(V, K, n_joints, verts, skin_indices, skin_weights,
                                                                      mats, out);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

static int skin_dense(bool backward, int batch, int V, int Vd, int J, const float *verts, const float *delta,
                      const float *weights, const int64_t *nn_ids, const float *A, const float *Bs, const float *Rh,
                      const float *Th, const float *dL_dout, float *out, cudaStream_t stream) {
    if (batch < 0 || V < 0 || Vd <= 0 || J <= 0) return D3GA_EINVAL;
    if (J > LBS_MAX_JOINTS) return D3GA_ELIMIT;
    if (batch == 0 || V == 0) return D3GA_OK;
    if (!weights || !nn_ids || !A || !Rh || !out) return D3GA_EINVAL;
    if (backward ? !dL_dout : (!verts || !Bs || !Th)) return D3GA_EINVAL;
    ProfScope prof(backward ? K_LBS_DENSE_BWD : K_LBS_DENSE_FWD, stream);
    const dim3 grid((V + LBS_THREADS - 1) / LBS_THREADS, batch);
    const size_t smem = (size_t)J * 12 * sizeof(float);
    if (backward)
        lbs_skin_dense_kernel<true><<<grid, LBS_THREADS, smem, stream>>>(V, Vd, J, verts, delta, weights, nn_ids, A, Bs, Rh,
                                                                         Th, dL_dout, out);
    else
        lbs_skin_dense_kernel<false><<<grid, LBS_THREADS, smem, stream>>>(V, Vd, J, verts, delta, weights, nn_ids, A, Bs, Rh,
                                                                          Th, dL_dout, out);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_lbs_skin_dense_forward(int batch, int V, int Vd, int J, const float *verts, const float *delta,
                                           const float *weights, const int64_t *nn_ids, const float *A, const float *Bs,
                                           const float *Rh, const float *Th, float *out, d3ga_stream_t stream) {
    return skin_dense(false, batch, V, Vd, J, verts, delta, weights, nn_ids, A, Bs, Rh, Th, nullptr, out,
                      (cudaStream_t)stream);
}

extern "C" int d3ga_lbs_skin_dense_backward(int batch, int V, int Vd, int J, const float *weights, const int64_t *nn_ids,
                                            const float *A, const float *Rh, const float *dL_dout, float *dL_dverts,
                                            d3ga_stream_t stream) {
    return skin_dense(true, batch, V, Vd, J, nullptr, nullptr, weights, nn_ids, A, nullptr, Rh, nullptr, dL_dout, dL_dverts,
                      (cudaStream_t)stream);
}
