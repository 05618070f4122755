"""ctypes binding of libd3ga_b200.so (the C-ABI declared in include/d3ga_b200.h).

There is no fallback: if the CUDA library is missing this module raises, and every op that needs it fails loudly.
PyTorch is used by the callers only for device memory and streams; nothing in the library links against it.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("D3GA_LIB") or os.path.join(_HERE, "libd3ga_b200.so")   # D3GA_LIB: development builds

# every symbol include/d3ga_b200.h declares (tests/test_abi.py checks the header against this list)
EXPORTS = [
    "d3ga_version", "d3ga_error_string", "d3ga_last_cuda_error", "d3ga_launch_count",
    "d3ga_profile_enable", "d3ga_profile_kernel_count", "d3ga_profile_kernel_name", "d3ga_profile_collect",
    "d3ga_raster_workspace_bytes", "d3ga_raster_forward_geometry", "d3ga_raster_forward_render",
    "d3ga_raster_forward_blend", "d3ga_raster_binning_path", "d3ga_set_views_in_flight", "d3ga_mark_visible", "d3ga_raster_sorted_keys",
    "d3ga_raster_backward_scratch_bytes", "d3ga_raster_backward", "d3ga_raster_backward_noalloc", "d3ga_raster_layout",
    "d3ga_raster_forward_render_pair", "d3ga_raster_backward_pair",
    "d3ga_deform_forward", "d3ga_deform_backward", "d3ga_deform_backward_ordered", "d3ga_deform_forward_fused",
    "d3ga_deform_backward_fused", "d3ga_deform_backward_fused_sparse", "d3ga_deform_backward_scratch_bytes",
    "d3ga_compute_bary",
    "d3ga_fem_energy_forward", "d3ga_fem_energy_backward", "d3ga_ssim_l1_forward", "d3ga_ssim_l1_backward",
    "d3ga_lbs_skeleton_forward", "d3ga_lbs_states_to_matrix", "d3ga_lbs_skin_forward", "d3ga_lbs_skin_backward", "d3ga_lbs_unskin_forward",
    "d3ga_lbs_skin_dense_forward", "d3ga_lbs_skin_dense_backward",
]

LAYOUT_NAMES = [
    "geom_header", "geom_depth", "geom_xy", "geom_conic_opacity", "geom_rgb", "geom_cov3D", "geom_tiles",
    "geom_offsets", "geom_clamped", "bin_keys0", "bin_keys1", "bin_vals0", "bin_vals1", "img_ranges", "img_final_T",
    "img_n_contrib",
]

_lib = None


class D3GAError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the library (once).  Raises D3GAError when it has not been built -- never falls back to anything."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise D3GAError(
            f"{LIB_PATH} is missing: build it with `python -m d3ga_b200.build` (nvcc, sm_100a). "
            "d3ga_b200 has no CPU or PyTorch fallback."
        )
    lib = C.CDLL(LIB_PATH)
    vp, i32, u64, f32 = C.c_void_p, C.c_int, C.c_uint64, C.c_float
    lib.d3ga_version.restype = C.c_int
    lib.d3ga_error_string.restype = C.c_char_p
    lib.d3ga_error_string.argtypes = [C.c_int]
    lib.d3ga_last_cuda_error.restype = C.c_int
    lib.d3ga_launch_count.restype = C.c_uint64
    lib.d3ga_profile_enable.argtypes = [C.c_int]
    lib.d3ga_profile_kernel_name.restype = C.c_char_p
    lib.d3ga_profile_kernel_name.argtypes = [C.c_int]
    lib.d3ga_profile_collect.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    lib.d3ga_raster_workspace_bytes.argtypes = [i32, i32, i32, u64, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
    lib.d3ga_raster_layout.argtypes = [i32, i32, i32, u64, C.POINTER(u64)]
    lib.d3ga_raster_binning_path.argtypes = [i32, i32, i32, u64]
    lib.d3ga_set_views_in_flight.argtypes = [i32]
    lib.d3ga_mark_visible.argtypes = [i32, vp, vp, vp, vp, vp]
    lib.d3ga_set_views_in_flight.restype = None
    lib.d3ga_raster_sorted_keys.argtypes = [i32, i32, i32, vp, vp, u64, vp, vp, vp]
    lib.d3ga_raster_forward_geometry.argtypes = (
        [i32, i32, i32, vp, vp, f32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, f32, i32, i32, vp, vp, u64, vp, vp, vp]
    )
    lib.d3ga_raster_forward_render.argtypes = [i32, i32, i32, vp, vp, vp, vp, u64, vp, vp, vp, vp]
    lib.d3ga_raster_forward_blend.argtypes = [i32, i32, i32, vp, vp, vp, vp, u64, vp, vp, vp, vp, vp]
    lib.d3ga_raster_backward_scratch_bytes.restype = C.c_uint64
    lib.d3ga_raster_backward_scratch_bytes.argtypes = [i32]
    lib.d3ga_raster_backward.argtypes = (
        [i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp, f32, f32, i32, vp, vp, vp, u64, vp,
         vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    )
    lib.d3ga_raster_forward_render_pair.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, u64, vp, vp, vp, vp]
    lib.d3ga_raster_backward_pair.argtypes = (
        [i32] * 5 + [vp] * 6 + [f32] + [vp] * 6 + [f32, f32, i32] + [vp] * 3 + [u64] + [vp] * 13
    )
    lib.d3ga_raster_backward_noalloc.argtypes = (
        [i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp, f32, f32, i32, vp, vp, vp, u64, vp,
         vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    )
    lib.d3ga_deform_forward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.d3ga_deform_backward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.d3ga_deform_backward_ordered.argtypes = [i32, i32, i32] + [vp] * 15
    lib.d3ga_deform_forward_fused.argtypes = [i32, i32, i32] + [vp] * 8 + [i32] + [vp] * 7
    lib.d3ga_deform_backward_fused.argtypes = [i32, i32, i32] + [vp] * 8 + [i32] + [vp] * 11 + [i32, i32, vp]
    lib.d3ga_deform_backward_fused_sparse.argtypes = [i32, i32, i32] + [vp] * 8 + [i32] + [vp] * 10 + [i32, i32, vp, vp]
    lib.d3ga_deform_backward_scratch_bytes.argtypes = [i32]
    lib.d3ga_deform_backward_scratch_bytes.restype = u64
    lib.d3ga_fem_energy_forward.argtypes = [i32, i32, vp, vp, vp, vp, vp]
    lib.d3ga_fem_energy_backward.argtypes = [i32, i32, vp, vp, vp, vp, vp, vp]
    lib.d3ga_ssim_l1_forward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp]
    lib.d3ga_ssim_l1_backward.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp]
    lib.d3ga_compute_bary.argtypes = [i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    i64 = C.c_int64
    lib.d3ga_lbs_skeleton_forward.argtypes = [i32, i32, i32, i32, vp, vp, i64, i32] + [vp] * 11
    lib.d3ga_lbs_states_to_matrix.argtypes = [i32, i32, vp, vp, vp, vp]
    lib.d3ga_lbs_skin_forward.argtypes = [i32, i32, i32, i32, vp, i64, vp, i64, f32, vp, vp, vp, vp, vp, i32, vp, f32, vp,
                                          vp, vp, vp]
    lib.d3ga_lbs_skin_backward.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, vp, i32, vp, f32, vp, f32, vp, vp]
    lib.d3ga_lbs_unskin_forward.argtypes = [i32, i32, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.d3ga_lbs_skin_dense_forward.argtypes = [i32, i32, i32, i32] + [vp] * 10
    lib.d3ga_lbs_skin_dense_backward.argtypes = [i32, i32, i32, i32] + [vp] * 7
    for name in EXPORTS:
        getattr(lib, name)   # AttributeError if the build is stale
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        lib = load()
        msg = lib.d3ga_error_string(rc).decode()
        extra = f" (cudaError {lib.d3ga_last_cuda_error()})" if rc == -2 else ""
        raise D3GAError(f"{what}: {msg}{extra}")


def ptr(t):
    """Device pointer of a tensor, or NULL for None / empty."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def workspace_bytes(P: int, W: int, H: int, capacity: int):
    lib = load()
    g, b, i = C.c_uint64(), C.c_uint64(), C.c_uint64()
    check(lib.d3ga_raster_workspace_bytes(P, W, H, capacity, C.byref(g), C.byref(b), C.byref(i)), "workspace_bytes")
    return g.value, b.value, i.value


def raster_layout(P: int, W: int, H: int, capacity: int) -> dict:
    lib = load()
    out = (C.c_uint64 * len(LAYOUT_NAMES))()
    check(lib.d3ga_raster_layout(P, W, H, capacity, out), "raster_layout")
    return dict(zip(LAYOUT_NAMES, [int(v) for v in out]))


_graph_launches = 0


def note_graph_launches(n: int) -> None:
    """Kernels of this library that a CUDA-graph replay launched (the C counter only sees launches made through the
    entry points, i.e. at capture time)."""
    global _graph_launches
    _graph_launches += int(n)


def launch_count() -> int:
    """Kernels launched by this library: through its entry points plus those replayed from captured graphs."""
    return int(load().d3ga_launch_count()) + _graph_launches


def profile_enable(on: bool) -> None:
    load().d3ga_profile_enable(1 if on else 0)


def profile_collect() -> dict:
    """{kernel name: (total ms, launches)} since the last collect."""
    lib = load()
    n = lib.d3ga_profile_kernel_count()
    ms = (C.c_double * n)()
    cnt = (C.c_uint64 * n)()
    check(lib.d3ga_profile_collect(ms, cnt), "profile_collect")
    return {lib.d3ga_profile_kernel_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}


# ---- cheap device / stream plumbing (torch.cuda.current_stream() and torch.cuda.device() cost ~20 us each) ------------
import torch as _torch


def raw_stream(device) -> int:
    """cudaStream_t of torch's current stream on `device`, as an integer."""
    return _torch._C._cuda_getCurrentRawStream(device.index if device.index is not None else _torch.cuda.current_device())


class device_guard:
    """Makes `device` current for the duration of a call (the C-ABI runs on the current device); no-op when it is."""

    __slots__ = ("idx", "prev")

    def __init__(self, device):
        self.idx = device.index if device.index is not None else -1
        self.prev = -1

    def __enter__(self):
        if self.idx >= 0:
            cur = _torch._C._cuda_getDevice()
            if cur != self.idx:
                self.prev = cur
                _torch._C._cuda_setDevice(self.idx)
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            _torch._C._cuda_setDevice(self.prev)
        return False
