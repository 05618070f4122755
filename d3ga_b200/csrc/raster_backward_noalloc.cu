// raster_backward_noalloc.cu -- d3ga_raster_backward_noalloc(): the rasterizer backward for hosts that must not
// allocate.
//
// d3ga_raster_backward() needs 48 bytes of scratch per Gaussian (the packed gradient accumulator that the blend
// backward fills and the per-Gaussian chain rule reads) and expects the caller to bring it.  Python callers get it
// from torch's caching allocator for free.  Some hosts have nothing comparable: a plain C/C++ playback or fitting
// loop that drives an avatar frame after frame has only cudaMalloc/cudaFree, which synchronise the whole device and
// can fail late once the process has committed its memory; deployments that hand out every byte of device memory at
// start-up have no allocator left at all by the time the loop runs.  For those callers this file keeps the scratch
// in the library image itself: a file-scope __device__ array of a fixed, modest size.  Every GPU that loads the
// library carries those D3GA_NOALLOC_MAX_GAUSSIANS * 48 bytes (12 MiB) until the process exits.
//
// The array's name is only a host-side handle; the device address is looked up with cudaGetSymbolAddress() the
// first time the entry point runs on a device (with that device current) and cached per device.  That first call
// should therefore be made before the stream is put into capture mode.
#include <mutex>

#include "common.cuh"

namespace d3ga {

constexpr int NOALLOC_MAX_GAUSSIANS = D3GA_NOALLOC_MAX_GAUSSIANS;
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) float g_noalloc_grad_acc[(size_t)NOALLOC_MAX_GAUSSIANS * GRAD_ACC_FLOATS];

static std::mutex g_noalloc_mu;
static float *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the static scratch on the CURRENT device (looked up once per device).
static int noalloc_scratch(float **out) {
    int dev = -1;
    D3GA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return D3GA_EINVAL;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        D3GA_CUDA(cudaGetSymbolAddress(&p, g_noalloc_grad_acc));
        g_noalloc_ptr[dev] = static_cast<float *>(p);
    }
    *out = g_noalloc_ptr[dev];
    return D3GA_OK;
}

bool profiling_enabled();   // api.cu

}  // namespace d3ga

using namespace d3ga;

extern "C" int d3ga_raster_backward_noalloc(
    int P, int sh_degree, int sh_coeffs, int W, int H, const float *background, const float *means3D,
    const float *colors_precomp, const float *opacities, const float *scales, float scale_modifier,
    const float *rotations, const float *cov3D_precomp, const float *shs, const float *viewmatrix,
    const float *projmatrix, const float *campos, float tanfovx, float tanfovy, int antialiasing,
    const int32_t *radii, const void *geom, const void *binning, uint64_t capacity, const void *image,
    const float *dL_dout_color, const float *dL_dout_invdepth, float *dL_dmeans2D, float *dL_dopacity,
    float *dL_dcolors, float *dL_dmeans3D, float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
    d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P > NOALLOC_MAX_GAUSSIANS) return D3GA_ELIMIT;   // the static scratch holds at most this many Gaussians
    if (profiling_enabled()) return D3GA_ENOTREADY;      // the timing hooks create CUDA events, which allocates
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !image || !background || !dL_dout_color || !viewmatrix ||
        !projmatrix || !campos)
        return D3GA_EINVAL;
    if (P == 0) return D3GA_OK;
    if (!means3D || !opacities || !radii || !dL_dmeans3D || !dL_dopacity) return D3GA_EINVAL;
    if ((cov3D_precomp != nullptr) == (scales != nullptr && rotations != nullptr)) return D3GA_EINVAL;
    if ((colors_precomp != nullptr) == (shs != nullptr)) return D3GA_EINVAL;
    if (shs && !dL_dsh) return D3GA_EINVAL;
    if (scales && (!dL_dscales || !dL_drotations)) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;

    float *acc = nullptr;
    int rc = noalloc_scratch(&acc);
    if (rc) return rc;
    // the array still holds the previous call's sums: the accumulator must start at zero
    D3GA_CUDA(cudaMemsetAsync(acc, 0, (size_t)P * GRAD_ACC_FLOATS * sizeof(float), stream));

    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    const ImgLayout IL = img_layout(P, W, H, capacity);
    const float *colors = colors_precomp ? colors_precomp : reinterpret_cast<const float *>((const uint8_t *)geom + GL.rgb);
    rc = launch_blend_backward(W, H, background, colors, (const uint8_t *)geom, GL, (const uint8_t *)binning, BL,
                               (const uint8_t *)image, IL, dL_dout_color, dL_dout_invdepth, acc, stream);
    if (rc) return rc;
    return launch_preprocess_backward(P, sh_degree, sh_coeffs, W, H, means3D, opacities, scales, scale_modifier,
                                      rotations, cov3D_precomp, shs, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                                      antialiasing, radii, (const uint8_t *)geom, GL, acc, dL_dout_invdepth != nullptr,
                                      dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
                                      dL_drotations, stream);
}
