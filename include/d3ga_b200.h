/*
 * d3ga_b200.h -- C-ABI of libd3ga_b200.so: the B200 (sm_100a) render hot path of D3GA.
 *
 * Plain pointers and sizes only; no torch types.  All pointers are DEVICE pointers unless a parameter says
 * "host".  Every entry point:
 *   - returns 0 on success or a negative D3GA_E* code (d3ga_error_string() gives the text);
 *   - enqueues its work on `stream` (a cudaStream_t passed as void*) and returns without synchronising;
 *   - allocates nothing: the caller owns outputs and workspaces (sizes from the *_workspace_bytes queries);
 *   - runs on the CURRENT device: the caller makes the buffers' device current first (the reference's own
 *     extensions never did -- sampler_points.cpp:191 takes the current stream without a device guard).
 *
 * Each declaration cites the reference interface it replaces.  Citations are relative to the D3GA tree.
 * The rasterizer's native source is a third-party dependency that is absent from that tree
 * (graphdeco-inria/diff-gaussian-rasterization @ 9c5c2028, .SUBMODULES.json:16-22), so those entry points cite
 * the reference's call site, which fixes names, argument meaning and layouts.
 */
#ifndef D3GA_B200_H
#define D3GA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define D3GA_OK 0
#define D3GA_EINVAL (-1)    /* bad argument (null pointer, negative size, inconsistent option set) */
#define D3GA_ECUDA (-2)     /* a CUDA runtime call failed; d3ga_last_cuda_error() has the cudaError_t */
#define D3GA_ETOOBIG (-3)   /* size beyond what the 32-bit duplicate indices / look-back words can hold */
#define D3GA_ENOTREADY (-4) /* call not possible in the current state */
#define D3GA_ELIMIT (-5)    /* size beyond a fixed capacity (d3ga_raster_backward_noalloc, d3ga_lbs_skeleton_forward) */

typedef void *d3ga_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------------------------------------ */
int d3ga_version(void);
const char *d3ga_error_string(int code);
int d3ga_last_cuda_error(void);
/* Number of kernels this library has launched in this process (the `gpu_launches` figure of bench.py). */
uint64_t d3ga_launch_count(void);

/* Measurement only (no counterpart in the reference; used by bench.py): when enabled, every kernel launch is
 * bracketed by a pair of CUDA events on its stream.  d3ga_profile_collect() waits for the recorded events, adds up
 * milliseconds and launches per kernel id (arrays of d3ga_profile_kernel_count() entries) and clears the record. */
void d3ga_profile_enable(int on);
int d3ga_profile_kernel_count(void);
const char *d3ga_profile_kernel_name(int id);
int d3ga_profile_collect(double *ms, uint64_t *launches);

/* ---- rasterizer -------------------------------------------------------------
 * Replaces: diff_gaussian_rasterization._C.rasterize_gaussians / rasterize_gaussians_backward as reached from
 * renderer.py:130-141 (GaussianRasterizer.__call__) with the settings of renderer.py:79-93.
 *
 * Layouts (all fp32 row-major as the reference passes them): means3D [P,3]; cov3D_precomp [P,6] packed
 * (xx,xy,xz,yy,yz,zz) (utils/general_utils.py:24-34) OR scales [P,3] + rotations [P,4] (r,x,y,z);
 * opacities [P]; colors_precomp [P,3] OR shs [P,sh_coeffs,3]; viewmatrix / projmatrix [16] in the row-vector
 * convention of lib/cameras.py:68-75; campos [3]; background [3]; out_color [3,H,W] planar; out_invdepth [H,W].
 *
 * Three workspaces, mirroring the upstream geomBuffer / binningBuffer / imgBuffer that the autograd function
 * keeps alive between forward and backward:
 *   geom    : per-Gaussian SoA (depth, xy, conic+opacity, rgb, cov3D, tiles_touched, offsets) + header + scan state
 *   binning : (tile|depth) keys and Gaussian indices, double-buffered, + radix-sort state; sized by `capacity`
 *   image   : per-tile ranges, per-pixel final transmittance and contributor count
 * `capacity` is the number of (tile, Gaussian) duplicates the binning workspace can hold.  The true count
 * ("num_rendered") is only known on the device; it is written to num_rendered_host (2 x uint32: count, overflow
 * flag) by an async copy at the end of the geometry stage.  If it exceeds `capacity` the later stages do nothing
 * and the caller re-runs BOTH stages with a larger capacity.  This replaces upstream's mid-pipeline host
 * synchronisation: the caller may enqueue the render stage first and check the count afterwards.
 */
int d3ga_raster_workspace_bytes(int P, int W, int H, uint64_t capacity,
                                uint64_t *geom_bytes, uint64_t *binning_bytes, uint64_t *image_bytes);

/* Replaces diff_gaussian_rasterization._C.mark_visible (GaussianRasterizer.markVisible upstream; D3GA's renderer.py never
 * calls it, it is part of the module's surface): present[i] = 1 when means3D[i] lies in front of the near plane
 * (z_view > 0.2), with the arithmetic the preprocess culls with.  present: uint8 [P] (a torch.bool tensor's storage).
 * projmatrix is accepted for signature parity and not read. */
int d3ga_mark_visible(int P, const float *means3D, const float *viewmatrix, const float *projmatrix, uint8_t *present,
                      d3ga_stream_t stream);

/* Sizing hint (no counterpart in the reference): how many independent views the caller keeps in flight on separate
 * streams (d3ga_b200.training.PipelinedViews does; default 1).  The direct tile scatter sizes its rows so that one view's
 * scatter occupies about 1/n of the GPU for n times as long -- alone that is slower, with n views in flight it leaves the
 * other views' kernels room beside it (measured at n = 4: +6 % views/s).  Results do not depend on it.  Process-wide, and
 * read by stage 1 (d3ga_raster_forward_geometry) ONLY: stage 1 files the row plan it chose under the geometry workspace
 * it ran on, and stage 2 / d3ga_raster_sorted_keys take the plan from there (D3GA_ENOTREADY when no stage 1 has run on
 * that workspace with the same sizes), so changing the hint between the two stages of a view -- from another thread,
 * say -- cannot make them disagree.  d3ga_b200.training sets it only while it captures a view's CUDA graph (the launch
 * configurations are baked into the graph) and restores 1, so eager calls keep the single-view sizing. */
void d3ga_set_views_in_flight(int n);

/* Which binning strategy the two forward stages will use for these sizes (measurement and tests; no counterpart in the
 * reference): 1 = direct tile scatter (the default where it applies: the N Gaussians are sorted by depth, stage 1 only
 * COUNTS the duplicates per tile, stage 2 stores every duplicate once, straight into its final slot of the sorted list --
 * no unsorted duplicate list exists after stage 1); 0 = a radix sort of the duplicate list that stage 1 emitted
 * (D3GA_SORT_MODE=depthfirst|global, tile grids beyond 49152 tiles, or a capacity too small to hold the
 * scatter's count rows).  Either way the sorted list, the ranges and num_rendered are bit-identical. */
int d3ga_raster_binning_path(int P, int W, int H, uint64_t capacity);

/* The sorted (tile << 32 | depth bits) key list of the last forward -- upstream's binningBuffer keys after
 * cub::DeviceRadixSort::SortPairs -- written to keys_out [num_rendered] (device memory, caller-owned).  Call after
 * d3ga_raster_forward_render.  The radix paths copy the list they sorted; the direct tile scatter never stores keys (the
 * blend reads Gaussian ids and ranges only, and 8 scattered bytes per duplicate cost more than the rest of the scatter),
 * so there the list is rebuilt from what defines it: slot i of tile t's range holds (t << 32) | bits(depth[id[i]]).
 * Used by the parity tests (bit-exact against the oracle's sorted keys) and for inspection. */
int d3ga_raster_sorted_keys(int P, int W, int H, const void *geom, const void *binning, uint64_t capacity,
                            const void *image, uint64_t *keys_out, d3ga_stream_t stream);

/* Stage 1 = preprocess (R1) + scan (R2) + key duplication (R3); with the direct tile scatter: preprocess, depth sort of
 * the Gaussians and the per-tile counts (which give num_rendered).
 * Exactly one of {cov3D_precomp} / {scales, rotations} and one of {colors_precomp} / {shs} is non-null.
 * radii [P] int32 is an output (the second element of the tuple GaussianRasterizer returns).
 * num_rendered_host: pinned HOST memory, 2 x uint32, may be null. */
int d3ga_raster_forward_geometry(int P, int sh_degree, int sh_coeffs,
                                 const float *means3D, const float *scales, float scale_modifier,
                                 const float *rotations, const float *opacities, const float *shs,
                                 const float *cov3D_precomp, const float *colors_precomp,
                                 const float *viewmatrix, const float *projmatrix, const float *campos,
                                 int W, int H, float tanfovx, float tanfovy, int prefiltered, int antialiasing,
                                 void *geom, void *binning, uint64_t capacity,
                                 int32_t *radii, uint32_t *num_rendered_host, d3ga_stream_t stream);

/* Stage 2 = sort (R4) + tile ranges (R5) + alpha blending (R6); with the direct tile scatter R2-R5 are the offsets from
 * stage 1's per-tile counts and one store of every duplicate into its final slot.  out_invdepth may be null. */
int d3ga_raster_forward_render(int P, int W, int H, const float *background, const float *colors_precomp,
                               void *geom, void *binning, uint64_t capacity, void *image,
                               float *out_color, float *out_invdepth, d3ga_stream_t stream);

/* Stage 2 for a PAIR of images of the same Gaussians from the same camera (the RGB and the
 * silhouette render of a D3GA training frame, models/trainer.py:102-110) in ONE walk of the tile lists: out_color is what
 * d3ga_raster_forward_render(background, colors_precomp) gives and out_color_b what a second render with
 * (background_b, colors_b [P,3]) on the same geometry would give -- same alphas, same transmittance, same contributor
 * set, so the two images are bit-identical to the two separate renders'.  No inverse depth in this form.
 * d3ga_raster_backward_pair is its backward: dL_dout_color and dL_dout_color_b [3,H,W] both flow into dL_dmeans3D,
 * dL_dmeans2D, dL_dopacity, dL_dcov3D / dL_dscales + dL_drotations; dL_dcolors (or dL_dsh) is the FIRST image's colour
 * gradient; colors_b receives none (in D3GA it is a constant, models/cage_net.py:243). */
int d3ga_raster_forward_render_pair(int P, int W, int H, const float *background, const float *colors_precomp,
                                    const float *background_b, const float *colors_b, void *geom, void *binning,
                                    uint64_t capacity, void *image, float *out_color, float *out_color_b,
                                    d3ga_stream_t stream);

/* Alpha blending (R6) alone, on the geometry and binning state an earlier d3ga_raster_forward_geometry +
 * d3ga_raster_forward_render pair left in `geom` / `binning` -- for a second render of the SAME Gaussians from the SAME
 * camera with other colours / background.  D3GA does exactly that every training frame: the RGB and the silhouette
 * render of models/trainer.py:102-110 share means, covariances and opacities, so R1-R5 need to run once
 * (SURVEY.md 8(f) rank 1).  `tile_src_image` is the image workspace of the render that did the binning (its per-tile
 * ranges and launch order are copied into `image`, which must be a different buffer).  The two renders then have
 * separate image workspaces and may be differentiated independently with d3ga_raster_backward. */
int d3ga_raster_forward_blend(int P, int W, int H, const float *background, const float *colors_precomp, void *geom,
                              const void *binning, uint64_t capacity, const void *tile_src_image, void *image,
                              float *out_color, float *out_invdepth, d3ga_stream_t stream);

/* Backward = blend backward (R7) + preprocess backward (R8).
 * Inputs are the forward's inputs and workspaces plus dL_dout_color [3,H,W] (and optionally dL_dout_invdepth
 * [H,W]).  Outputs (written, not accumulated): dL_dmeans3D [P,3], dL_dmeans2D [P,3] (may be null: not wanted), dL_dopacity [P],
 * dL_dcolors [P,3] (when colors_precomp was given) or dL_dsh [P,sh_coeffs,3], dL_dcov3D [P,6] (when
 * cov3D_precomp was given) or dL_dscales [P,3] + dL_drotations [P,4].
 * grad_scratch: d3ga_raster_backward_scratch_bytes(P) bytes, contents undefined on entry. */
uint64_t d3ga_raster_backward_scratch_bytes(int P);
int d3ga_raster_backward(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                         const float *means3D, const float *colors_precomp, const float *opacities,
                         const float *scales, float scale_modifier, const float *rotations,
                         const float *cov3D_precomp, const float *shs,
                         const float *viewmatrix, const float *projmatrix, const float *campos,
                         float tanfovx, float tanfovy, int antialiasing, const int32_t *radii,
                         const void *geom, const void *binning, uint64_t capacity, const void *image,
                         const float *dL_dout_color, const float *dL_dout_invdepth,
                         void *grad_scratch,
                         float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors, float *dL_dmeans3D,
                         float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                         d3ga_stream_t stream);

int d3ga_raster_backward_pair(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                              const float *background_b, const float *means3D, const float *colors_precomp,
                              const float *opacities, const float *scales, float scale_modifier, const float *rotations,
                              const float *cov3D_precomp, const float *shs, const float *viewmatrix,
                              const float *projmatrix, const float *campos, float tanfovx, float tanfovy,
                              int antialiasing, const int32_t *radii, const void *geom, const void *binning,
                              uint64_t capacity, const void *image, const float *dL_dout_color,
                              const float *dL_dout_color_b, void *grad_scratch, float *dL_dmeans2D, float *dL_dopacity,
                              float *dL_dcolors, float *dL_dmeans3D, float *dL_dcov3D, float *dL_dsh, float *dL_dscales,
                              float *dL_drotations, d3ga_stream_t stream);

/* The same backward for hosts that cannot allocate device memory (no counterpart in the reference, whose Python
 * glue always has torch's allocator): a C/C++ playback or fitting loop without a caching allocator, where
 * cudaMalloc/cudaFree would stall the whole device and may fail once memory is committed, or a deployment that hands
 * out all device memory at start-up.  Identical to d3ga_raster_backward() except that there is no grad_scratch
 * argument: the 48-byte-per-Gaussian accumulator lives in a static __device__ array inside the library, which every
 * GPU that loads the library holds (D3GA_NOALLOC_MAX_GAUSSIANS * 48 B = 12 MiB) until the process exits.
 *   - P > D3GA_NOALLOC_MAX_GAUSSIANS returns D3GA_ELIMIT; nothing is launched.
 *   - It calls no allocating routine; while d3ga_profile_enable(1) is in effect it returns D3GA_ENOTREADY, because
 *     the timing hooks create CUDA events.
 *   - The first call on a device looks the array's device address up (with that device current) and caches it; make
 *     that call before any stream capture begins.
 *   - Every call uses the same array and clears the part it needs first.  Two calls whose GPU work can overlap
 *     (different streams, or different threads on one device) race on it: keep them on one stream, or serialise. */
#define D3GA_NOALLOC_MAX_GAUSSIANS 262144
int d3ga_raster_backward_noalloc(int P, int sh_degree, int sh_coeffs, int W, int H, const float *background,
                                 const float *means3D, const float *colors_precomp, const float *opacities,
                                 const float *scales, float scale_modifier, const float *rotations,
                                 const float *cov3D_precomp, const float *shs,
                                 const float *viewmatrix, const float *projmatrix, const float *campos,
                                 float tanfovx, float tanfovy, int antialiasing, const int32_t *radii,
                                 const void *geom, const void *binning, uint64_t capacity, const void *image,
                                 const float *dL_dout_color, const float *dL_dout_invdepth,
                                 float *dL_dmeans2D, float *dL_dopacity, float *dL_dcolors, float *dL_dmeans3D,
                                 float *dL_dcov3D, float *dL_dsh, float *dL_dscales, float *dL_drotations,
                                 d3ga_stream_t stream);

/* Byte offsets of the sections inside the three workspaces, for parity tests that compare key lists, sorted
 * indices and ranges with the oracle.  out[] receives D3GA_LAYOUT_COUNT values, indexed by the enum. */
enum {
    D3GA_L_GEOM_HEADER = 0, /* uint32[64]: [0] num_rendered, [1] overflow, [3] final sort buffer (0/1)      */
    D3GA_L_GEOM_DEPTH,      /* float  [P]                                                                  */
    D3GA_L_GEOM_XY,         /* float2 [P]                                                                  */
    D3GA_L_GEOM_CONIC_OPACITY, /* float4 [P]                                                               */
    D3GA_L_GEOM_RGB,        /* float  [P,3] (SH path; the colors_precomp path reads the caller's tensor)   */
    D3GA_L_GEOM_COV3D,      /* float  [P,6] (scales/rotations path)                                        */
    D3GA_L_GEOM_TILES,      /* uint32 [P] tiles touched                                                    */
    D3GA_L_GEOM_OFFSETS,    /* uint32 [P] inclusive prefix sum of tiles touched, in emission order (radix paths)*/
    D3GA_L_GEOM_CLAMPED,    /* uint8  [P,3]                                                                */
    D3GA_L_BIN_KEYS0,       /* uint64 [capacity]  radix paths (d3ga_raster_binning_path() == 0): unsorted keys after
                             * stage 1 (emission order: by depth with D3GA_SORT_MODE=depthfirst, by Gaussian index
                             * with =global); the header's final_sel word says which buffer holds the sorted list.
                             * Direct tile scatter: not written -- use d3ga_raster_sorted_keys()              */
    D3GA_L_BIN_KEYS1,       /* uint64 [capacity]                                                           */
    D3GA_L_BIN_VALS0,       /* uint32 [capacity]  radix paths: unsorted Gaussian indices after stage 1; direct tile
                             * scatter: the sorted indices after stage 2                                     */
    D3GA_L_BIN_VALS1,       /* uint32 [capacity]                                                           */
    D3GA_L_IMG_RANGES,      /* uint2  [tiles]                                                              */
    D3GA_L_IMG_FINAL_T,     /* float  [H*W]                                                                */
    D3GA_L_IMG_N_CONTRIB,   /* uint32 [H*W]                                                                */
    D3GA_LAYOUT_COUNT
};
int d3ga_raster_layout(int P, int W, int H, uint64_t capacity, uint64_t *out);

/* ---- cage -> Gaussian deformation -------------------------------------------
 * Replaces the torch op chain of models/cage_net.py:218-230 (+ lib/cage.py:339-342, tetra_sampler/tetra.py:74-80,
 * utils/general_utils.py:24-34,58-90, models/cage_net.py:161-164) by one kernel each way.
 * tetpoints [V,3] f32; tets [T,4] int64; tetra_id [N] int64; barys [N,4] f32 (not renormalised);
 * canonical_gradient [N,3,3] f32; scales [N,3] f32 (post-exp); rotation [N,4] f32 (r,x,y,z; normalised inside).
 * Outputs: means3D [N,3], cov3D [N,6].  A negative tetra_id wraps to T + id like torch indexing does. */
int d3ga_deform_forward(int N, int V, int T, const float *tetpoints, const int64_t *tets, const int64_t *tetra_id,
                        const float *barys, const float *canonical_gradient, const float *scales,
                        const float *rotation, float *means3D, float *cov3D, d3ga_stream_t stream);
/* dL_dtetpoints [V,3] is ACCUMULATED into (the caller zeroes it; several garments / views may add to it);
 * dL_dbarys [N,4], dL_dscales [N,3], dL_drotation [N,4] are written.  Any output pointer may be null. */
int d3ga_deform_backward(int N, int V, int T, const float *tetpoints, const int64_t *tets, const int64_t *tetra_id,
                         const float *barys, const float *canonical_gradient, const float *scales,
                         const float *rotation, const float *dL_dmeans3D, const float *dL_dcov3D,
                         float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales, float *dL_drotation,
                         d3ga_stream_t stream);

/* Same backward with the Gaussians visited in the order `order` [N] int32 gives (a permutation of 0..N-1, normally
 * argsort(tetra_id), which the caller computes once -- tetra_id is a constant buffer, lib/cage.py:331-337).  Any
 * permutation is valid; when it groups equal tetra_id together the cage-vertex gradients of a tet are summed inside
 * the warp and one lane per tet issues the atomics (~N/T times fewer than d3ga_deform_backward). */
int d3ga_deform_backward_ordered(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                 const int64_t *tetra_id, const float *barys, const float *canonical_gradient,
                                 const float *scales, const float *rotation, const float *dL_dmeans3D,
                                 const float *dL_dcov3D, float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales,
                                 float *dL_drotation, const int32_t *order, d3ga_stream_t stream);

/* The deformation block of models/cage_net.py:213-230 in ONE launch each way: the plain entry points above plus
 *   canon_barys = barys + delta_bary                       (cage_net.py:213; delta_bary [N,4] or null)
 *   scales      = exp(scaling + delta_scale)               (cage_net.py:214; log_scales != 0: `scales` and delta_scale
 *                                                           [N,3] are pre-activation; delta_scale may be null)
 *   rotation    = F.normalize(rotation + delta_rot)        (cage_net.py:215; delta_rot [N,4] or null; normalised once
 *                                                           inside -- build_rotation normalises again in the reference,
 *                                                           so the result is the same; the zero vector is NaN in both)
 *   fm_energy   = fem_energy(tetpoints)  per tet           (cage_net.py:225, lib/cage.py:349-361; dn_inv [T,3,3] + energy
 *                                                           [T] both given or both null) -- run by extra CTAs of the
 *                                                           same launch, it gathers the same tetpoints[tets]
 * Backward: a base tensor and its delta have the same gradient, written ONCE (dL_dbarys = dL/dbarys = dL/ddelta_bary,
 * dL_dscales is the gradient of the PRE-activation scales when log_scales != 0, dL_drotation of the un-normalised sum);
 * dL_denergy [T] or null adds the energy's gradient into dL_dtetpoints (which the caller zeroes, as above).
 * `order` as in d3ga_deform_backward_ordered, or null.  rows_tet_sorted != 0 (with order == null) tells the kernel that
 * tetra_id is non-decreasing along the rows as they lie in memory: the Gaussians are then read in natural order --
 * coalesced -- and a tet's cage-vertex gradients are still summed inside the warp before the atomics.  It is a hint:
 * results are right for any tetra_id, only the number of atomics depends on it.  dL_dtetpoints_pitch = floats per row of
 * dL_dtetpoints: 3 (the tensor's own layout) or 4 (a padded [V,4] buffer, 16-byte aligned: a vertex's three components
 * then go out as one vector reduction instead of three scalar atomics; the fourth float stays 0). */
int d3ga_deform_forward_fused(int N, int V, int T, const float *tetpoints, const int64_t *tets, const int64_t *tetra_id,
                              const float *barys, const float *delta_bary, const float *canonical_gradient,
                              const float *scales, const float *delta_scale, int log_scales, const float *rotation,
                              const float *delta_rot, const float *dn_inv, float *energy, float *means3D,
                              float *cov3D, d3ga_stream_t stream);
int d3ga_deform_backward_fused(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                               const int64_t *tetra_id, const float *barys, const float *delta_bary,
                               const float *canonical_gradient, const float *scales, const float *delta_scale,
                               int log_scales, const float *rotation, const float *delta_rot, const float *dn_inv,
                               const float *dL_dmeans3D, const float *dL_dcov3D, const float *dL_denergy,
                               float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales, float *dL_drotation,
                               const int32_t *order, int rows_tet_sorted, int dL_dtetpoints_pitch,
                               d3ga_stream_t stream);

/* The same backward for the gradients a RENDERED view hands back: most Gaussians were never blended (everything behind
 * the surface that saturated the pixels) and arrive with dL_dmeans3D = dL_dcov3D = 0 exactly -- 86-93 % of a 1 M-Gaussian
 * avatar per camera.  A first kernel streams the incoming gradients, writes those Gaussians' zero gradients itself and
 * lists the others; the backward proper then runs over the list alone.  Results are identical to
 * d3ga_deform_backward_fused (order = null); `scratch` = d3ga_deform_backward_scratch_bytes(N) bytes of device memory,
 * contents undefined on entry.  (With gradients that are dense -- every Gaussian non-zero -- it is the slower form.) */
uint64_t d3ga_deform_backward_scratch_bytes(int N);
int d3ga_deform_backward_fused_sparse(int N, int V, int T, const float *tetpoints, const int64_t *tets,
                                      const int64_t *tetra_id, const float *barys, const float *delta_bary,
                                      const float *canonical_gradient, const float *scales, const float *delta_scale,
                                      int log_scales, const float *rotation, const float *delta_rot, const float *dn_inv,
                                      const float *dL_dmeans3D, const float *dL_dcov3D, const float *dL_denergy,
                                      float *dL_dtetpoints, float *dL_dbarys, float *dL_dscales, float *dL_drotation,
                                      int rows_tet_sorted, int dL_dtetpoints_pitch, void *scratch, d3ga_stream_t stream);

/* ---- cage FEM energy ------------------------------------------------------------
 * Replaces CageBase.fem_energy (lib/cage.py:349-361, called at models/cage_net.py:225) and its autograd backward:
 * per tet, F = G(tetpoints[tets]) @ dn_inv, energy = 0.5 (det F - 1)^2 + 0.5 (tr(F^T F) - 3).
 * tetpoints [V,3] f32; tets [T,4] int64; dn_inv [T,3,3] f32 (inverse rest-pose edge matrices, lib/cage.py:312-313);
 * energy [T] f32.  The backward ACCUMULATES into dL_dtetpoints [V,3] (caller zeroes it). */
int d3ga_fem_energy_forward(int T, int V, const float *tetpoints, const int64_t *tets, const float *dn_inv,
                            float *energy, d3ga_stream_t stream);
int d3ga_fem_energy_backward(int T, int V, const float *tetpoints, const int64_t *tets, const float *dn_inv,
                             const float *dL_denergy, float *dL_dtetpoints, d3ga_stream_t stream);

/* ---- fused L1 + SSIM image loss -------------------------------------------------
 * Replaces utils/loss_utils.py:29-30 (l1_loss) and :59-90 (ssim, window 11, sigma 1.5, zero padding) with their
 * autograd backward, as combined at train.py:190-193.  img1 (prediction), img2 (target): [C,H,W] f32.
 * Forward: sums[0] = sum of the SSIM map, sums[1] = sum of |img1 - img2| (device, 2 floats, overwritten); the three
 * derivative maps [C,H,W] are optional (all or none) and only needed for the backward.
 * Backward: dL_dimg1 = u * (g_ssim * dSSIMsum/dimg1 + g_l1 * sign(img1 - img2))  (written); u = *upstream, the scalar
 * dL/dloss in DEVICE memory (autograd's grad_output: no host read-back, so the loss can sit inside a captured graph),
 * or 1 when upstream is null. */
int d3ga_ssim_l1_forward(int C, int H, int W, const float *img1, const float *img2, float *sums, float *dm_dmu1,
                         float *dm_dE11, float *dm_dE12, d3ga_stream_t stream);
int d3ga_ssim_l1_backward(int C, int H, int W, const float *img1, const float *img2, const float *dm_dmu1,
                          const float *dm_dE11, const float *dm_dE12, float g_ssim, float g_l1, const float *upstream,
                          float *dL_dimg1, d3ga_stream_t stream);

/* ---- cage-vertex linear blend skinning -------------------------------------------
 * The step that produces `tetpoints` right before the deformation.  Quaternions are (x, y, z, w); joint matrices
 * are row-major [3][4] (rotation*scale | translation), the layout of states_to_matrix (lbsmodel/body_model.py:350-387).
 *
 * d3ga_lbs_skeleton_forward replaces ParameterTransform.forward (body_model.py:39-47), solve_skeleton_state
 * (:311-347) and states_to_matrix (:350-387): one kernel, one CTA per pose.  The raw model parameters are
 * (pose [batch, n_pose] | scales [n_scale]); scales_stride = floats between poses (0 = one scale vector for all,
 * the lbs_scale.expand of body_model.py:670); scales may be null (all zero).  with_zero_scale_copy != 0 appends
 * `batch` more solves of the same poses with every scale parameter zero -- the solve
 * compute_root_rigid_transform (body_model.py:176-190) needs -- so states / mats then hold 2*batch poses.
 * The parameter transform [7*n_joints, n_pose+n_scale] is passed as CSR (row_ptr [7*NJ+1], col_idx, vals) plus
 * offsets [7*NJ]; joint_offset [NJ,3]; joint_rotation [NJ,4] pre-rotations; joint_parents [NJ] int32 (negative =
 * root; any index order works); bind_state [NJ,8] (required for mats).  Outputs (either may be null):
 * states [*,NJ,8] = (t, q, scale), mats [*,NJ,3,4].  At most 1024 joints (D3GA_ELIMIT beyond).
 * d3ga_lbs_states_to_matrix is states_to_matrix alone, for callers that hold skeleton states. */
int d3ga_lbs_skeleton_forward(int batch, int n_joints, int n_pose, int n_scale, const float *pose, const float *scales,
                              int64_t scales_stride, int with_zero_scale_copy, const int32_t *row_ptr,
                              const int32_t *col_idx, const float *vals, const float *offsets,
                              const float *joint_offset, const float *joint_rotation, const int32_t *joint_parents,
                              const float *bind_state, float *states, float *mats, d3ga_stream_t stream);
int d3ga_lbs_states_to_matrix(int batch, int n_joints, const float *bind_state, const float *states, float *mats,
                              d3ga_stream_t stream);

/* d3ga_lbs_skin_forward replaces LinearBlendSkinning.skinning (body_model.py:208-234), LBSModule.pose's global
 * scaling (:669-675) and the tail of Blueman.get (lib/blueman.py:101-116,139-158; CageBase.transform lib/cage.py:54-58):
 *   v   = delta ? (verts / delta_unit + delta) * delta_unit : verts                         [V,3] per pose
 *   g   = global_scaling * sum_k skin_weights[v,k] * (mats[skin_indices[v,k]] @ [v;1])      K influences
 *   out = Rinv (g / unit_div) + tinv + shift,   [Rinv|tinv] = inverse([R|t/unit_div](root_mats[root_joint]) @ pre_mul)
 * verts_stride / delta_stride: floats between poses (0 = one array shared by all poses).  global_scaling [3],
 * root_mats [batch,NJ,3,4] (the zero-scale solve, body_model.py:176-190), pre_mul [4,4] row-major, shift [3] and
 * rt_out [batch,4,4] (the inverse transform, blueman.py:148) may each be null (identity / none).
 * d3ga_lbs_skin_backward writes dL_dverts [batch,V,3] = grad_scale * d out / d v ^T dL_dout (pass delta_unit as
 * grad_scale to differentiate with respect to delta).  Gradients to the pose parameters are not provided: the
 * reference never optimises them (models/cage_net.py:186-189 feed dataset motion). */
int d3ga_lbs_skin_forward(int batch, int V, int K, int n_joints, const float *verts, int64_t verts_stride,
                          const float *delta, int64_t delta_stride, float delta_unit, const int64_t *skin_indices,
                          const float *skin_weights, const float *mats, const float *global_scaling,
                          const float *root_mats, int root_joint, const float *pre_mul, float unit_div,
                          const float *shift, float *out, float *rt_out, d3ga_stream_t stream);
int d3ga_lbs_skin_backward(int batch, int V, int K, int n_joints, const int64_t *skin_indices,
                           const float *skin_weights, const float *mats, const float *global_scaling,
                           const float *root_mats, int root_joint, const float *pre_mul, float unit_div,
                           const float *dL_dout, float grad_scale, float *dL_dverts, d3ga_stream_t stream);

/* LinearBlendSkinning.unskinning (body_model.py:248-283, init only): blend each vertex's matrices, invert, apply.
 * verts, out [batch,V,3]. */
int d3ga_lbs_unskin_forward(int batch, int V, int K, int n_joints, const float *verts, const int64_t *skin_indices,
                            const float *skin_weights, const float *mats, float *out, d3ga_stream_t stream);

/* Smplman.deform (lib/smplman.py:155-171): out = Rodrigues(Rh) (T [v + delta + Bs[nn]; 1]) + Th with
 * T = sum_j weights[nn, j] A[j], nn = nn_ids[v].  verts, delta (nullable) [V,3]; weights [Vd,J]; nn_ids [V] int64;
 * A [batch,J,4,4]; Bs [batch,Vd,3]; Rh, Th [batch,3]; out [batch,V,3].  Backward: dL_dverts [batch,V,3]
 * (= gradient to delta as well). */
int d3ga_lbs_skin_dense_forward(int batch, int V, int Vd, int J, const float *verts, const float *delta,
                                const float *weights, const int64_t *nn_ids, const float *A, const float *Bs,
                                const float *Rh, const float *Th, float *out, d3ga_stream_t stream);
int d3ga_lbs_skin_dense_backward(int batch, int V, int Vd, int J, const float *weights, const int64_t *nn_ids,
                                 const float *A, const float *Rh, const float *dL_dout, float *dL_dverts,
                                 d3ga_stream_t stream);

/* ---- barycentric embedding -------------------------------------------------------
 * Replaces tetra_sampler.compute_bary (submodules/tetra-sampler/tetra_sampler/barys.py:9-33): the nearest-face
 * query of bvh_distance_queries.BVH (bvh_cuda_op.cu:1065-1269) followed by barycentric_kernel
 * (src/sampler_points_kernel.cu:279-313).  points [N,3]; tetras [T,4,3]; triangles [F,3,3]; tri_to_tetra [F,2]
 * int32 (-1 = no neighbour).  Outputs: barys [N,4] f32, tetra_id [N] int32 (-1 on miss, barys 0),
 * active [N] int32 (1 / -1), closest_face [N] int32 (may be null). */
int d3ga_compute_bary(int N, int T, int F, const float *points, const float *tetras, const float *triangles,
                      const int32_t *tri_to_tetra, float *barys, int32_t *tetra_id, int32_t *active,
                      int32_t *closest_face, d3ga_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* D3GA_B200_H */
