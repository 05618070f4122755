// api.cu -- the extern "C" surface of libd3ga_b200.so declared in include/d3ga_b200.h (rasterizer part; the
// deformation and embedding entry points live beside their kernels in deform.cu / bary.cu).
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace d3ga {

static std::atomic<uint64_t> g_launches{0};
static std::atomic<int> g_last_cuda_error{0};

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }
int set_cuda_error(cudaError_t e) {
    g_last_cuda_error.store((int)e);
    (void)cudaGetLastError();   // clear the sticky-free error state
    return D3GA_ECUDA;
}

// ---- per-kernel event timing -------------------------------------------------------------------------------------
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
struct ProfRec { int id; cudaEvent_t a, b; };
static std::vector<ProfRec> g_prof_recs;
static const char *const g_kernel_names[K_COUNT] = {
    "geometry_kernel", "sort_histogram_kernel", "sort_pass_kernel", "tile_ranges_kernel", "blend_forward_kernel",
    "blend_backward_kernel", "preprocess_backward_kernel", "deform_forward_kernel", "deform_backward_kernel",
    "compute_bary_kernel", "preprocess_kernel", "emit_sorted_kernel", "depth_plan_kernel",
    "depth_sort_pass_kernel", "fem_energy_forward_kernel", "fem_energy_backward_kernel",
    "ssim_forward_kernel", "ssim_backward_kernel", "lbs_skeleton_kernel", "lbs_skin_forward_kernel",
    "lbs_skin_backward_kernel", "lbs_unskin_kernel", "lbs_skin_dense_forward_kernel", "lbs_skin_dense_backward_kernel",
    "rect_grid_kernel", "tile_plan_kernel", "direct_count_kernel", "direct_segsum_kernel", "direct_plan_kernel",
    "direct_base_kernel", "direct_scatter_kernel", "pack_features_kernel", "deform_select_kernel"};

bool profiling_enabled() { return g_prof_on.load(std::memory_order_relaxed) != 0; }

ProfScope::ProfScope(int id_, cudaStream_t s) : id(id_), stream(s), slot(nullptr) {
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    ProfRec *r = new ProfRec;
    r->id = id;
    if (cudaEventCreate(&r->a) != cudaSuccess || cudaEventCreate(&r->b) != cudaSuccess) { delete r; return; }
    cudaEventRecord(r->a, stream);
    slot = r;
}
ProfScope::~ProfScope() {
    if (!slot) return;
    ProfRec *r = (ProfRec *)slot;
    cudaEventRecord(r->b, stream);
    std::lock_guard<std::mutex> lk(g_prof_mu);
    g_prof_recs.push_back(*r);
    delete r;
}

int device_sm_count() {
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev >= 0 && dev < 64 && cached[dev].load(std::memory_order_relaxed) > 0) return cached[dev].load(std::memory_order_relaxed);
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
    if (dev >= 0 && dev < 64) cached[dev].store(n, std::memory_order_relaxed);
    return n;
}

int sort_mode() {
    static const int mode = [] {
        const char *v = getenv("D3GA_SORT_MODE");
        if (v && strcmp(v, "global") == 0) return (int)SORT_GLOBAL;
        if (v && strcmp(v, "depthfirst") == 0) return (int)SORT_DEPTH_FIRST;
        return (int)SORT_DIRECT;
    }();
    return mode;
}

// upstream getHigherMsb: number of bits needed for the tile id
static int higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return (int)msb;
}

}  // namespace d3ga

using namespace d3ga;

extern "C" {

int d3ga_version(void) { return 100; }

const char *d3ga_error_string(int code) {
    switch (code) {
        case D3GA_OK: return "ok";
        case D3GA_EINVAL: return "invalid argument";
        case D3GA_ECUDA: return "CUDA runtime error (see d3ga_last_cuda_error)";
        case D3GA_ETOOBIG: return "size exceeds the 2^31-1 duplicate limit";
        case D3GA_ENOTREADY: return "call not possible in the current state (staged call out of order, or profiling on)";
        case D3GA_ELIMIT: return "size beyond a fixed capacity (the fixed scratch of the allocation-free entry point; 1024 joints of the skeleton kernel)";
        default: return "unknown error";
    }
}

int d3ga_last_cuda_error(void) { return g_last_cuda_error.load(); }

void d3ga_profile_enable(int on) { g_prof_on.store(on ? 1 : 0); }
int d3ga_profile_kernel_count(void) { return K_COUNT; }
const char *d3ga_profile_kernel_name(int id) { return id >= 0 && id < K_COUNT ? g_kernel_names[id] : ""; }
int d3ga_profile_collect(double *ms, uint64_t *launches) {
    std::vector<ProfRec> recs;
    {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        recs.swap(g_prof_recs);
    }
    for (int k = 0; k < K_COUNT; k++) { if (ms) ms[k] = 0.0; if (launches) launches[k] = 0; }
    for (auto &r : recs) {
        float t = 0.f;
        D3GA_CUDA(cudaEventSynchronize(r.b));
        D3GA_CUDA(cudaEventElapsedTime(&t, r.a, r.b));
        if (ms) ms[r.id] += (double)t;
        if (launches) launches[r.id] += 1;
        cudaEventDestroy(r.a);
        cudaEventDestroy(r.b);
    }
    return D3GA_OK;
}
uint64_t d3ga_launch_count(void) { return g_launches.load(); }

int d3ga_raster_workspace_bytes(int P, int W, int H, uint64_t capacity, uint64_t *geom_bytes, uint64_t *binning_bytes,
                                uint64_t *image_bytes) {
    if (P < 0 || W <= 0 || H <= 0) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    if (geom_bytes) *geom_bytes = geom_layout(P).total;
    if (binning_bytes) *binning_bytes = bin_layout(capacity).total;
    if (image_bytes) *image_bytes = img_layout(P, W, H, capacity).total;
    return D3GA_OK;
}

void d3ga_set_views_in_flight(int n) { set_views_in_flight(n); }

int d3ga_mark_visible(int P, const float *means3D, const float *viewmatrix, const float *projmatrix, uint8_t *present,
                      d3ga_stream_t stream) {
    (void)projmatrix;   // upstream passes it to in_frustum, which only uses it for a value it never tests
    if (P < 0 || !viewmatrix || (P > 0 && (!means3D || !present))) return D3GA_EINVAL;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

int d3ga_raster_binning_path(int P, int W, int H, uint64_t capacity) {
    if (P < 0 || W <= 0 || H <= 0) return D3GA_EINVAL;
    return direct_plan(P, (W + TILE - 1) / TILE, (H + TILE - 1) / TILE, capacity).ok ? 1 : 0;
}

int d3ga_raster_sorted_keys(int P, int W, int H, const void *geom, const void *binning, uint64_t capacity, const void *image,
                            uint64_t *keys_out, d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !image || !keys_out) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    const ImgLayout IL = img_layout(P, W, H, capacity);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    DirectPlan direct;
    if (!recall_direct_plan(geom, P, gx, gy, capacity, &direct)) return D3GA_ENOTREADY;
    if (direct.ok)
        return launch_sorted_keys(gx * gy, (const uint8_t *)geom, GL, (const uint8_t *)binning, BL, (const uint8_t *)image, IL,
                                  keys_out, stream);
    // radix paths: the list is in one of the two key buffers; the header says which (device-side selection, no host sync)
    return launch_copy_sorted_keys((const uint8_t *)geom, GL, (const uint8_t *)binning, BL, capacity, keys_out, stream);
}

int d3ga_raster_layout(int P, int W, int H, uint64_t capacity, uint64_t *out) {
    if (P < 0 || W <= 0 || H <= 0 || !out) return D3GA_EINVAL;
    const GeomLayout G = geom_layout(P);
    const BinLayout B = bin_layout(capacity);
    const ImgLayout I = img_layout(P, W, H, capacity);
    out[D3GA_L_GEOM_HEADER] = G.header; out[D3GA_L_GEOM_DEPTH] = G.depth; out[D3GA_L_GEOM_XY] = G.xy;
    out[D3GA_L_GEOM_CONIC_OPACITY] = G.conic_opacity; out[D3GA_L_GEOM_RGB] = G.rgb; out[D3GA_L_GEOM_COV3D] = G.cov3D;
    out[D3GA_L_GEOM_TILES] = G.tiles; out[D3GA_L_GEOM_OFFSETS] = G.offsets; out[D3GA_L_GEOM_CLAMPED] = G.clamped;
    out[D3GA_L_BIN_KEYS0] = B.keys0; out[D3GA_L_BIN_KEYS1] = B.keys1; out[D3GA_L_BIN_VALS0] = B.vals0;
    out[D3GA_L_BIN_VALS1] = B.vals1;
    out[D3GA_L_IMG_RANGES] = I.ranges; out[D3GA_L_IMG_FINAL_T] = I.final_T; out[D3GA_L_IMG_N_CONTRIB] = I.n_contrib;
    return D3GA_OK;
}

int d3ga_raster_forward_geometry(int P, int sh_degree, int sh_coeffs, const float *means3D, const float *scales,
                                 float scale_modifier, const float *rotations, const float *opacities,
                                 const float *shs, const float *cov3D_precomp, const float *colors_precomp,
                                 const float *viewmatrix, const float *projmatrix, const float *campos, int W, int H,
                                 float tanfovx, float tanfovy, int prefiltered, int antialiasing, void *geom,
                                 void *binning, uint64_t capacity, int32_t *radii, uint32_t *num_rendered_host,
                                 d3ga_stream_t stream_) {
    (void)prefiltered;   // upstream only uses it to trap on culled points in debug builds
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !viewmatrix || !projmatrix || !campos) return D3GA_EINVAL;
    if (P > 0 && (!means3D || !opacities || !radii)) return D3GA_EINVAL;
    if ((cov3D_precomp != nullptr) == (scales != nullptr && rotations != nullptr) && P > 0) return D3GA_EINVAL;
    if ((colors_precomp != nullptr) == (shs != nullptr) && P > 0) return D3GA_EINVAL;
    if (shs && (sh_degree < 0 || sh_degree > 3 || sh_coeffs < (sh_degree + 1) * (sh_degree + 1))) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    int rc = launch_geometry(P, sh_degree, sh_coeffs, means3D, scales, scale_modifier, rotations, opacities, shs,
                             cov3D_precomp, colors_precomp, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                             antialiasing, (uint8_t *)geom, GL, (uint8_t *)binning, BL, capacity, radii, stream);
    if (rc) return rc;
    if (num_rendered_host)
        D3GA_CUDA(cudaMemcpyAsync(num_rendered_host, (uint8_t *)geom + GL.header, 8, cudaMemcpyDeviceToHost, stream));
    return D3GA_OK;
}

static int forward_render_impl(int P, int W, int H, const float *background, const float *colors_precomp, void *geom,
                               void *binning, uint64_t capacity, void *image, float *out_color, float *out_invdepth,
                               const float *background_b, const float *colors_b, float *out_color_b,
                               d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !image || !background || !out_color) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    const ImgLayout IL = img_layout(P, W, H, capacity);
    const int num_tiles = ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    const int sort_bits = 32 + higher_msb((uint32_t)num_tiles);
    int rc = launch_sort_and_ranges(P, (uint8_t *)geom, GL, (W + TILE - 1) / TILE, (H + TILE - 1) / TILE, (uint8_t *)binning,
                                    BL, capacity, sort_bits, (uint8_t *)image, IL, num_tiles, stream);
    if (rc) return rc;
    const float *colors = colors_precomp ? colors_precomp : reinterpret_cast<const float *>((uint8_t *)geom + GL.rgb);
    return launch_blend_forward(P, W, H, background, colors, (const uint8_t *)geom, GL, (const uint8_t *)binning, BL,
                                (uint8_t *)image, IL, out_color, out_invdepth, stream, background_b, colors_b, out_color_b);
}

int d3ga_raster_forward_render(int P, int W, int H, const float *background, const float *colors_precomp, void *geom,
                               void *binning, uint64_t capacity, void *image, float *out_color, float *out_invdepth,
                               d3ga_stream_t stream) {
    return forward_render_impl(P, W, H, background, colors_precomp, geom, binning, capacity, image, out_color,
                               out_invdepth, nullptr, nullptr, nullptr, stream);
}

int d3ga_raster_forward_render_pair(int P, int W, int H, const float *background, const float *colors_precomp,
                                    const float *background_b, const float *colors_b, void *geom, void *binning,
                                    uint64_t capacity, void *image, float *out_color, float *out_color_b,
                                    d3ga_stream_t stream) {
    if (!background_b || !colors_b || !out_color_b) return D3GA_EINVAL;
    return forward_render_impl(P, W, H, background, colors_precomp, geom, binning, capacity, image, out_color, nullptr,
                               background_b, colors_b, out_color_b, stream);
}

int d3ga_raster_forward_blend(int P, int W, int H, const float *background, const float *colors_precomp, void *geom,
                              const void *binning, uint64_t capacity, const void *tile_src_image, void *image,
                              float *out_color, float *out_invdepth, d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !image || !tile_src_image || !background || !out_color)
        return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    const ImgLayout IL = img_layout(P, W, H, capacity);
    const uint64_t tiles = (uint64_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    // per-tile ranges and the forward launch order come from the render that did the binning
    D3GA_CUDA(cudaMemcpyAsync((uint8_t *)image + IL.ranges, (const uint8_t *)tile_src_image + IL.ranges, 8 * tiles,
                              cudaMemcpyDeviceToDevice, stream));
    D3GA_CUDA(cudaMemcpyAsync((uint8_t *)image + IL.order_fwd, (const uint8_t *)tile_src_image + IL.order_fwd,
                              4 * tiles, cudaMemcpyDeviceToDevice, stream));
    RasterHeader *hdr = reinterpret_cast<RasterHeader *>((uint8_t *)geom + GL.header);
    D3GA_CUDA(cudaMemsetAsync(&hdr->blend_done, 0, sizeof(uint32_t), stream));
    const float *colors = colors_precomp ? colors_precomp : reinterpret_cast<const float *>((uint8_t *)geom + GL.rgb);
    return launch_blend_forward(P, W, H, background, colors, (const uint8_t *)geom, GL, (const uint8_t *)binning, BL,
                                (uint8_t *)image, IL, out_color, out_invdepth, stream);
}

uint64_t d3ga_raster_backward_scratch_bytes(int P) { return align256((uint64_t)(P > 0 ? P : 1) * GRAD_ACC_FLOATS * 4); }

static int backward_impl(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                         const float *means3D, const float *colors_precomp, const float *opacities,
                         const float *scales, float scale_modifier, const float *rotations,
                         const float *cov3D_precomp, const float *shs, const float *viewmatrix,
                         const float *projmatrix, const float *campos, float tanfovx, float tanfovy, int antialiasing,
                         const int32_t *radii, const void *geom, const void *binning, uint64_t capacity,
                         const void *image, const float *dL_dout_color, const float *dL_dout_invdepth,
                         void *grad_scratch, float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors,
                         float *dL_dmeans3D, float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                         const float *background_b, const float *dL_dout_color_b, d3ga_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (P < 0 || W <= 0 || H <= 0 || !geom || !binning || !image || !background || !dL_dout_color || !grad_scratch ||
        !viewmatrix || !projmatrix || !campos)
        return D3GA_EINVAL;
    if (P == 0) return D3GA_OK;
    if (!means3D || !opacities || !radii || !dL_dmeans3D || !dL_dopacity) return D3GA_EINVAL;
    if ((cov3D_precomp != nullptr) == (scales != nullptr && rotations != nullptr)) return D3GA_EINVAL;
    if ((colors_precomp != nullptr) == (shs != nullptr)) return D3GA_EINVAL;
    if (shs && !dL_dsh) return D3GA_EINVAL;
    if (scales && (!dL_dscales || !dL_drotations)) return D3GA_EINVAL;
    if (capacity > MAX_DUPLICATES) return D3GA_ETOOBIG;
    const GeomLayout GL = geom_layout(P);
    const BinLayout BL = bin_layout(capacity);
    const ImgLayout IL = img_layout(P, W, H, capacity);
    float *acc = (float *)grad_scratch;
    D3GA_CUDA(cudaMemsetAsync(acc, 0, (size_t)P * GRAD_ACC_FLOATS * 4, stream));
    const float *colors = colors_precomp ? colors_precomp : reinterpret_cast<const float *>((const uint8_t *)geom + GL.rgb);
    int rc = launch_blend_backward(W, H, background, colors, (const uint8_t *)geom, GL, (const uint8_t *)binning, BL,
                                   (const uint8_t *)image, IL, dL_dout_color, dL_dout_invdepth, acc, stream, background_b,
                                   dL_dout_color_b);
    if (rc) return rc;
    return launch_preprocess_backward(P, sh_degree, sh_coeffs, W, H, means3D, opacities, scales, scale_modifier,
                                      rotations, cov3D_precomp, shs, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                                      antialiasing, radii, (const uint8_t *)geom, GL, acc, dL_dout_invdepth != nullptr,
                                      dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
                                      dL_drotations, stream);
}

int d3ga_raster_backward(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                         const float *means3D, const float *colors_precomp, const float *opacities,
                         const float *scales, float scale_modifier, const float *rotations,
                         const float *cov3D_precomp, const float *shs, const float *viewmatrix,
                         const float *projmatrix, const float *campos, float tanfovx, float tanfovy, int antialiasing,
                         const int32_t *radii, const void *geom, const void *binning, uint64_t capacity,
                         const void *image, const float *dL_dout_color, const float *dL_dout_invdepth,
                         void *grad_scratch, float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors,
                         float *dL_dmeans3D, float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                         d3ga_stream_t stream) {
    return backward_impl(P, sh_degree, sh_coeffs, W, H, background, means3D, colors_precomp, opacities, scales,
                         scale_modifier, rotations, cov3D_precomp, shs, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                         antialiasing, radii, geom, binning, capacity, image, dL_dout_color, dL_dout_invdepth, grad_scratch,
                         dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
                         nullptr, nullptr, stream);
}

int d3ga_raster_backward_pair(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                              const float *background_b, const float *means3D, const float *colors_precomp,
                              const float *opacities, const float *scales, float scale_modifier, const float *rotations,
                              const float *cov3D_precomp, const float *shs, const float *viewmatrix,
                              const float *projmatrix, const float *campos, float tanfovx, float tanfovy,
                              int antialiasing, const int32_t *radii, const void *geom, const void *binning,
                              uint64_t capacity, const void *image, const float *dL_dout_color,
                              const float *dL_dout_color_b, void *grad_scratch, float *dL_dmeans2D, float *dL_dopacity,
                              float *dL_dcolors, float *dL_dmeans3D, float *dL_dcov3D, float *dL_dsh, float *dL_dscales,
                              float *dL_drotations, d3ga_stream_t stream) {
    if (!background_b || !dL_dout_color_b) return D3GA_EINVAL;
    return backward_impl(P, sh_degree, sh_coeffs, W, H, background, means3D, colors_precomp, opacities, scales,
                         scale_modifier, rotations, cov3D_precomp, shs, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                         antialiasing, radii, geom, binning, capacity, image, dL_dout_color, nullptr, grad_scratch,
                         dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
                         background_b, dL_dout_color_b, stream);
}

}  // extern "C"
