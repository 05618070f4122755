"""Helpers of the -m gpu parity tests: drive the C-ABI stage by stage and pull the workspaces back as numpy."""
import numpy as np
import torch

from d3ga_b200 import _lib


def dev(a, dtype=torch.float32):
    return None if a is None else torch.as_tensor(np.ascontiguousarray(a)).to("cuda", dtype).contiguous()


def section(buf, off, count, dtype):
    nbytes = count * np.dtype(dtype).itemsize
    return buf[off:off + nbytes].cpu().numpy().view(dtype).copy()


def forward_stages(g, cam, W, H, bg, path="precomp", aa=False, capacity=None, invdepth=True, stop_after=None,
                   between_stages=None):
    """Runs d3ga_raster_forward_geometry (+ d3ga_raster_forward_render) and returns every intermediate as numpy."""
    lib = _lib.load()
    P = g["means"].shape[0]
    means, op = dev(g["means"]), dev(g["opacity"])
    sh = col = cov = sc = rot = None
    M = 0
    if path == "precomp":
        col, cov = dev(g["colors"]), dev(g["cov6"])
    else:
        sh, sc, rot = dev(g["shs"]), dev(g["scales"]), dev(g["rot"])
        M = sh.shape[1]
    view, proj, campos, bgt = dev(cam.viewmatrix), dev(cam.projmatrix), dev(cam.campos), dev(np.asarray(bg, np.float32))
    cap = capacity if capacity is not None else max(1024, 64 * P)
    gb, bb, ib = _lib.workspace_bytes(P, W, H, cap)
    geom = torch.zeros(gb, dtype=torch.uint8, device="cuda")
    binning = torch.zeros(bb, dtype=torch.uint8, device="cuda")
    image = torch.zeros(ib, dtype=torch.uint8, device="cuda")
    radii = torch.zeros(max(P, 1), dtype=torch.int32, device="cuda")
    host = torch.zeros(2, dtype=torch.int32).pin_memory()
    stream = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.d3ga_raster_forward_geometry(
        P, 3 if sh is not None else 0, M, _lib.ptr(means), _lib.ptr(sc), 1.0, _lib.ptr(rot), _lib.ptr(op), _lib.ptr(sh),
        _lib.ptr(cov), _lib.ptr(col), view.data_ptr(), proj.data_ptr(), campos.data_ptr(), W, H, cam.tanfovx,
        cam.tanfovy, 0, int(aa), geom.data_ptr(), binning.data_ptr(), cap, radii.data_ptr(), host.data_ptr(), stream),
        "forward_geometry")
    torch.cuda.synchronize()
    lay = _lib.raster_layout(P, W, H, cap)
    D, overflow = int(host[0]) & 0xFFFFFFFF, int(host[1])
    out = dict(P=P, num_rendered=D, overflow=overflow, capacity=cap, radii=radii[:P].cpu().numpy(), layout=lay)
    # 1 = direct tile scatter: stage 1 only counts, there is no unsorted duplicate list and no per-Gaussian offsets
    out["direct"] = int(lib.d3ga_raster_binning_path(P, W, H, cap)) == 1
    out["depths"] = section(geom, lay["geom_depth"], P, np.float32)
    out["xy"] = section(geom, lay["geom_xy"], 2 * P, np.float32).reshape(P, 2)
    out["conic_opacity"] = section(geom, lay["geom_conic_opacity"], 4 * P, np.float32).reshape(P, 4)
    out["tiles_touched"] = section(geom, lay["geom_tiles"], P, np.uint32)
    out["offsets"] = section(geom, lay["geom_offsets"], P, np.uint32)
    if sh is not None:
        out["rgb"] = section(geom, lay["geom_rgb"], 3 * P, np.float32).reshape(P, 3)
        out["clamped"] = section(geom, lay["geom_clamped"], 3 * P, np.uint8).reshape(P, 3)
        out["cov3D"] = section(geom, lay["geom_cov3D"], 6 * P, np.float32).reshape(P, 6)
    n = min(D, cap)
    out["keys_unsorted"] = section(binning, lay["bin_keys0"], n, np.uint64)
    out["vals_unsorted"] = section(binning, lay["bin_vals0"], n, np.uint32)
    out["tensors"] = dict(means=means, op=op, sh=sh, col=col, cov=cov, sc=sc, rot=rot, view=view, proj=proj,
                          campos=campos, bg=bgt, geom=geom, binning=binning, image=image, radii=radii, M=M)
    if stop_after == "geometry" or overflow:
        return out
    if between_stages is not None:
        between_stages()
    color = torch.zeros((3, H, W), dtype=torch.float32, device="cuda")
    inv = torch.zeros((1, H, W), dtype=torch.float32, device="cuda") if invdepth else None
    _lib.check(lib.d3ga_raster_forward_render(P, W, H, bgt.data_ptr(), _lib.ptr(col), geom.data_ptr(),
                                              binning.data_ptr(), cap, image.data_ptr(), color.data_ptr(),
                                              _lib.ptr(inv), stream), "forward_render")
    torch.cuda.synchronize()
    hdr = section(geom, lay["geom_header"], 64, np.uint32)
    out["header"] = hdr
    sel = int(hdr[3])
    keys = torch.zeros(max(D, 1), dtype=torch.int64, device="cuda")
    _lib.check(lib.d3ga_raster_sorted_keys(P, W, H, geom.data_ptr(), binning.data_ptr(), cap, image.data_ptr(),
                                           keys.data_ptr(), stream), "sorted_keys")
    torch.cuda.synchronize()
    out["keys"] = keys[:D].cpu().numpy().view(np.uint64).copy()
    out["point_list"] = section(binning, lay["bin_vals1" if sel else "bin_vals0"], D, np.uint32)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = section(image, lay["img_ranges"], 2 * tiles, np.uint32).reshape(tiles, 2)
    out["final_T"] = section(image, lay["img_final_T"], W * H, np.float32)
    out["n_contrib"] = section(image, lay["img_n_contrib"], W * H, np.uint32)
    out["color"] = color.cpu().numpy()
    out["invdepth"] = inv.cpu().numpy() if inv is not None else None
    return out


def backward(out, W, H, cam, dL_dcolor, dL_dinv=None, aa=False):
    lib = _lib.load()
    t = out["tensors"]
    P, M = out["P"], t["M"]
    gcol = dev(dL_dcolor)
    ginv = dev(dL_dinv) if dL_dinv is not None else None
    z = lambda *s: torch.full(s, float("nan"), dtype=torch.float32, device="cuda")  # noqa: E731
    scratch = torch.empty(int(lib.d3ga_raster_backward_scratch_bytes(P)), dtype=torch.uint8, device="cuda")
    res = dict(means2D=z(P, 3), opacity=z(P), means3D=z(P, 3))
    res["colors"] = z(P, 3) if t["col"] is not None else None
    res["cov3D"] = z(P, 6) if t["cov"] is not None else None
    res["sh"] = z(P, M, 3) if t["sh"] is not None else None
    res["scales"] = z(P, 3) if t["sc"] is not None else None
    res["rotations"] = z(P, 4) if t["rot"] is not None else None
    _lib.check(lib.d3ga_raster_backward(
        P, 3 if t["sh"] is not None else 0, M, W, H, t["bg"].data_ptr(), _lib.ptr(t["means"]), _lib.ptr(t["col"]),
        _lib.ptr(t["op"]), _lib.ptr(t["sc"]), 1.0, _lib.ptr(t["rot"]), _lib.ptr(t["cov"]), _lib.ptr(t["sh"]),
        t["view"].data_ptr(), t["proj"].data_ptr(), t["campos"].data_ptr(), cam.tanfovx, cam.tanfovy, int(aa),
        t["radii"].data_ptr(), t["geom"].data_ptr(), t["binning"].data_ptr(), out["capacity"], t["image"].data_ptr(),
        gcol.data_ptr(), _lib.ptr(ginv), scratch.data_ptr(), _lib.ptr(res["means2D"]), _lib.ptr(res["opacity"]),
        _lib.ptr(res["colors"]), _lib.ptr(res["means3D"]), _lib.ptr(res["cov3D"]), _lib.ptr(res["sh"]),
        _lib.ptr(res["scales"]), _lib.ptr(res["rotations"]), torch.cuda.current_stream().cuda_stream), "backward")
    torch.cuda.synchronize()
    return {k: (v.cpu().numpy() if v is not None else None) for k, v in res.items()}
