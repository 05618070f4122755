// ssim.cu -- SURVEY.md section 8(f) rank 3: fused L1 + SSIM image loss, forward and backward.
//
// Replaces utils/loss_utils.py:29-30 (l1_loss) and :59-90 (ssim / _ssim: five grouped 11x11 Gaussian convolutions and
// ~15 elementwise kernels over [3,H,W], plus their autograd backward) as used at train.py:190-193:
//     color = (1 - lambda_dssim) * l1_loss(pred, gt) + lambda_dssim * (1 - ssim(pred, gt))
// One kernel each way.  The 11-tap window (sigma 1.5, normalised in fp32 like gaussian() / create_window()) is applied
// separably from a shared-memory tile (16x16 outputs + 5-pixel halo); zero padding like F.conv2d(padding=5).
// Forward: per-pixel SSIM, block-reduced sums of SSIM and |pred - gt| (two fp32 atomics per CTA), and the three partial
// derivative maps dS/dmu1, dS/dE[x^2], dS/dE[xy] that the backward convolves.  Pure stencil, HBM-bound:
// algorithmic bytes per pixel-channel 8 read + 12 written forward, 20 read + 4 written backward.
#include "common.cuh"

namespace d3ga {
namespace {

constexpr int SS_T = 16, SS_R = 5, SS_W = SS_T + 2 * SS_R;   // 26
struct SsimWindow { float w[11]; };

__global__ void __launch_bounds__(SS_T * SS_T) ssim_forward_kernel(int H, int W, const float *__restrict__ img1,
                                                                  const float *__restrict__ img2, SsimWindow win,
                                                                  float *__restrict__ sums /* [2]: ssim, l1 */,
                                                                  float *__restrict__ dm_dmu1,
                                                                  float *__restrict__ dm_dE11,
                                                                  float *__restrict__ dm_dE12) {
    __shared__ float s1[SS_W][SS_W + 1], s2[SS_W][SS_W + 1];
    __shared__ float h[5][SS_W][SS_T + 1];
    __shared__ float s_red[2][8];
    const int tx = threadIdx.x % SS_T, ty = threadIdx.x / SS_T, tid = threadIdx.x;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t plane = (size_t)blockIdx.z * H * W;
    for (int i = tid; i < SS_W * SS_W; i += SS_T * SS_T) {
        const int r = i / SS_W, c = i % SS_W, y = y0 + r - SS_R, x = x0 + c - SS_R;
        const bool in = y >= 0 && y < H && x >= 0 && x < W;
        s1[r][c] = in ? img1[plane + (size_t)y * W + x] : 0.f;
        s2[r][c] = in ? img2[plane + (size_t)y * W + x] : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < SS_W * SS_T; i += SS_T * SS_T) {
        const int r = i / SS_T, c = i % SS_T;
        float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float p = s1[r][c + k], q = s2[r][c + k], wk = win.w[k];
            a += wk * p; b += wk * q; aa += wk * p * p; bb += wk * q * q; ab += wk * p * q;
        }
        h[0][r][c] = a; h[1][r][c] = b; h[2][r][c] = aa; h[3][r][c] = bb; h[4][r][c] = ab;
    }
    __syncthreads();
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float wk = win.w[k];
        mu1 += wk * h[0][ty + k][tx]; mu2 += wk * h[1][ty + k][tx];
        e11 += wk * h[2][ty + k][tx]; e22 += wk * h[3][ty + k][tx]; e12 += wk * h[4][ty + k][tx];
    }
    const int x = x0 + tx, y = y0 + ty;
    const bool inside = x < W && y < H;
    float ssim = 0.f, l1 = 0.f;
    if (inside) {
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
        const float s1sq = e11 - mu1_sq, s2sq = e22 - mu2_sq, s12 = e12 - mu12;
        const float A1 = 2.f * mu12 + C1, A2 = 2.f * s12 + C2, B1 = mu1_sq + mu2_sq + C1, B2 = s1sq + s2sq + C2;
        const float inv = 1.f / (B1 * B2);
        ssim = A1 * A2 * inv;
        const size_t o = plane + (size_t)y * W + x;
        if (dm_dmu1) {
            // S = A1 A2 / (B1 B2);  E11 = E[x^2], E12 = E[xy] are held fixed for the mu1 partial
            const float dS_dA1 = A2 * inv, dS_dA2 = A1 * inv, dS_dB1 = -ssim / B1, dS_dB2 = -ssim / B2;
            dm_dmu1[o] = dS_dA1 * 2.f * mu2 + dS_dB1 * 2.f * mu1 - dS_dA2 * 2.f * mu2 - dS_dB2 * 2.f * mu1;
            dm_dE11[o] = dS_dB2;
            dm_dE12[o] = 2.f * dS_dA2;
        }
        l1 = fabsf(s1[ty + SS_R][tx + SS_R] - s2[ty + SS_R][tx + SS_R]);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) {
        ssim += __shfl_xor_sync(0xffffffffu, ssim, d);
        l1 += __shfl_xor_sync(0xffffffffu, l1, d);
    }
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = ssim; s_red[1][tid >> 5] = l1; }
    __syncthreads();
    if (tid < 2) {
        float t = 0.f;
        for (int w = 0; w < 8; w++) t += s_red[tid][w];
        atomicAdd(sums + tid, t);
    }
}

// dL/dimg1 = g_ssim * [ conv(dm_dmu1) + 2 img1 conv(dm_dE11) + img2 conv(dm_dE12) ] + g_l1 * sign(img1 - img2)
__global__ void __launch_bounds__(SS_T * SS_T) ssim_backward_kernel(int H, int W, const float *__restrict__ img1,
                                                                   const float *__restrict__ img2, SsimWindow win,
                                                                   const float *__restrict__ dm_dmu1,
                                                                   const float *__restrict__ dm_dE11,
                                                                   const float *__restrict__ dm_dE12, float g_ssim,
                                                                   float g_l1, const float *__restrict__ upstream,
                                                                   float *__restrict__ dL_dimg1) {
    __shared__ float s[3][SS_W][SS_W + 1];
    __shared__ float h[3][SS_W][SS_T + 1];
    const int tx = threadIdx.x % SS_T, ty = threadIdx.x / SS_T, tid = threadIdx.x;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t plane = (size_t)blockIdx.z * H * W;
    for (int i = tid; i < SS_W * SS_W; i += SS_T * SS_T) {
        const int r = i / SS_W, c = i % SS_W, y = y0 + r - SS_R, x = x0 + c - SS_R;
        const bool in = y >= 0 && y < H && x >= 0 && x < W;
        const size_t o = plane + (size_t)y * W + x;
        s[0][r][c] = in ? dm_dmu1[o] : 0.f;
        s[1][r][c] = in ? dm_dE11[o] : 0.f;
        s[2][r][c] = in ? dm_dE12[o] : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < SS_W * SS_T; i += SS_T * SS_T) {
        const int r = i / SS_T, c = i % SS_T;
        float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float wk = win.w[k];
            a += wk * s[0][r][c + k]; b += wk * s[1][r][c + k]; d += wk * s[2][r][c + k];
        }
        h[0][r][c] = a; h[1][r][c] = b; h[2][r][c] = d;
    }
    __syncthreads();
    const int x = x0 + tx, y = y0 + ty;
    if (x >= W || y >= H) return;
    float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float wk = win.w[k];
        a += wk * h[0][ty + k][tx]; b += wk * h[1][ty + k][tx]; d += wk * h[2][ty + k][tx];
    }
    const size_t o = plane + (size_t)y * W + x;
    const float p = img1[o], q = img2[o];
    const float diff = p - q;
    const float sgn = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
    const float up = upstream ? __ldg(upstream) : 1.f;   // dL/dloss, a device scalar: no host read-back in autograd
    dL_dimg1[o] = up * (g_ssim * (a + 2.f * p * b + q * d) + g_l1 * sgn);
}

}  // namespace
}  // namespace d3ga

using namespace d3ga;

static SsimWindow make_window() {
    // utils/loss_utils.py:46-48: exp(-(x - 5)^2 / (2 * 1.5^2)) as Python floats, stored as float32, divided by their sum
    SsimWindow w;
    float g[11], sum = 0.f;
    for (int x = 0; x < 11; x++) { g[x] = (float)exp(-((double)(x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5)); sum += g[x]; }
    for (int x = 0; x < 11; x++) w.w[x] = g[x] / sum;
    return w;
}

extern "C" int d3ga_ssim_l1_forward(int C, int H, int W, const float *img1, const float *img2, float *sums,
                                    float *dm_dmu1, float *dm_dE11, float *dm_dE12, d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !sums) return D3GA_EINVAL;
    if ((dm_dmu1 != nullptr) != (dm_dE11 != nullptr) || (dm_dmu1 != nullptr) != (dm_dE12 != nullptr)) return D3GA_EINVAL;
    D3GA_CUDA(cudaMemsetAsync(sums, 0, 2 * sizeof(float), stream));
    const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C);
    ProfScope prof(K_SSIM_FWD, stream);
    ssim_forward_kernel<<<grid, SS_T * SS_T, 0, stream>>>(H, W, img1, img2, make_window(), sums, dm_dmu1, dm_dE11, dm_dE12);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}

extern "C" int d3ga_ssim_l1_backward(int C, int H, int W, const float *img1, const float *img2, const float *dm_dmu1,
                                     const float *dm_dE11, const float *dm_dE12, float g_ssim, float g_l1,
                                     const float *upstream, float *dL_dimg1, d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !dm_dmu1 || !dm_dE11 || !dm_dE12 || !dL_dimg1) return D3GA_EINVAL;
    const dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T, C);
    ProfScope prof(K_SSIM_BWD, stream);
    ssim_backward_kernel<<<grid, SS_T * SS_T, 0, stream>>>(H, W, img1, img2, make_window(), dm_dmu1, dm_dE11, dm_dE12,
                                                           g_ssim, g_l1, upstream, dL_dimg1);
    D3GA_LAUNCH_CHECK();
    return D3GA_OK;
}
