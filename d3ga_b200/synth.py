"""Synthetic avatars of the shapes BASELINE.json names; numpy only, seeded, fp32.

Nothing here is a model: it produces the *inputs of the hot path* -- a tetrahedral cage, Gaussians embedded in its
tets, a smooth deformation of the cage vertices and cameras in the reference's matrix convention
(lib/cameras.py:62-75, utils/graphics_utils.py:41-75: row-vector transforms, znear 0.01, zfar 100).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

# Kuhn split of the unit cube into 6 tets that all share the main diagonal (v0 -> v7); faces of neighbouring
# cubes match, so the lattice is a conforming tet mesh.
_KUHN = np.array(
    [[0, 1, 3, 7], [0, 3, 2, 7], [0, 2, 6, 7], [0, 6, 4, 7], [0, 4, 5, 7], [0, 5, 1, 7]], dtype=np.int64
)


def make_cage(nx: int, ny: int, nz: int, box=(0.6, 1.8, 0.4), origin=(-0.3, -0.9, -0.2)):
    """Axis-aligned lattice of nx*ny*nz cubes, each split into 6 tets.  Returns (points[V,3] f32, tets[T,4] i64)."""
    xs = np.linspace(0.0, box[0], nx + 1) + origin[0]
    ys = np.linspace(0.0, box[1], ny + 1) + origin[1]
    zs = np.linspace(0.0, box[2], nz + 1) + origin[2]
    gx, gy, gz = np.meshgrid(xs, ys, zs, indexing="ij")
    pts = np.stack([gx, gy, gz], -1).reshape(-1, 3).astype(np.float32)

    def vid(i, j, k):
        return (i * (ny + 1) + j) * (nz + 1) + k

    i, j, k = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    i, j, k = i.ravel(), j.ravel(), k.ravel()
    corners = np.stack(
        [vid(i + (c & 1), j + ((c >> 1) & 1), k + ((c >> 2) & 1)) for c in range(8)], axis=1
    )  # [cubes, 8]
    tets = corners[:, _KUHN].reshape(-1, 4).astype(np.int64)
    return pts, tets


def deform_points(x: np.ndarray, t=(0.01, 0.02, -0.015)) -> np.ndarray:
    """Smooth deformation with det J > 0: Rz(0.3 rad) x + 0.02 sin(2 pi x / 0.5) + t."""
    c, s = math.cos(0.3), math.sin(0.3)
    Rz = np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)
    y = x.astype(np.float64) @ Rz.T + 0.02 * np.sin(2 * math.pi * x.astype(np.float64) / 0.5) + np.asarray(t)
    return y.astype(np.float32)


def tet_gradient(X: np.ndarray) -> np.ndarray:
    """Edge matrix with (x3-x0, x2-x0, x1-x0) as COLUMNS (tetra_sampler/tetra.py:74-80)."""
    return np.stack([X[:, 3] - X[:, 0], X[:, 2] - X[:, 0], X[:, 1] - X[:, 0]], axis=2)


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: np.ndarray  # [4,4] f32, world_view_transform (transposed W2C)
    projmatrix: np.ndarray  # [4,4] f32, full_proj_transform
    campos: np.ndarray  # [3] f32


def camera_pose(W: int, H: int, azim_deg=0.0, elev_deg=0.0, dist=3.0, center=(0.0, 0.0, 0.0), focal_mul=1.2):
    """What a dataset hands the reference's Camera (lib/cameras.py:29-44): R (camera-to-world rotation, columns = camera
    axes, x right / y down / z forward), T (world-to-camera translation), FoVx, FoVy."""
    az, el = math.radians(azim_deg), math.radians(elev_deg)
    center = np.asarray(center, np.float64)
    eye = center + dist * np.array([math.sin(az) * math.cos(el), math.sin(el), math.cos(az) * math.cos(el)])
    f = center - eye
    f /= np.linalg.norm(f)
    up = np.array([0.0, 1.0, 0.0])
    right = np.cross(f, up)
    right /= np.linalg.norm(right)
    down = np.cross(f, right)
    R = np.stack([right, down, f], axis=1)
    T = -R.T @ eye
    focal = focal_mul * max(W, H)
    return R, T, 2 * math.atan(W / (2 * focal)), 2 * math.atan(H / (2 * focal))


def make_camera(W: int, H: int, azim_deg=0.0, elev_deg=0.0, dist=3.0, center=(0.0, 0.0, 0.0), focal_mul=1.2) -> Camera:
    """The matrices lib/cameras.py:62-75 builds from camera_pose(...) (pinned to the reference's own Camera class by
    tests/golden/camera_golden.npz)."""
    R, T, _, _ = camera_pose(W, H, azim_deg, elev_deg, dist, center, focal_mul)
    # getWorld2View2 (utils/graphics_utils.py:41-52)
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    w2c = np.float32(np.linalg.inv(np.linalg.inv(Rt)))
    view = np.ascontiguousarray(w2c.T)
    focal = focal_mul * max(W, H)
    fovx, fovy = 2 * math.atan(W / (2 * focal)), 2 * math.atan(H / (2 * focal))
    znear, zfar = 0.01, 100.0
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right_ = ty * znear, tx * znear
    Pm = np.zeros((4, 4), np.float32)  # getProjectionMatrix (utils/graphics_utils.py:55-75)
    Pm[0, 0] = 2.0 * znear / (2 * right_)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    proj = np.ascontiguousarray(Pm.T)
    full = (view @ proj).astype(np.float32)
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return Camera(W, H, math.tan(fovx * 0.5), math.tan(fovy * 0.5), view, np.ascontiguousarray(full), campos)


def camera_ring(W: int, H: int, n: int, elevations=(-15.0, 0.0, 15.0, 30.0)):
    """C5: n cameras = len(elevations) elevations x n/len azimuths."""
    per = max(1, n // len(elevations))
    cams = []
    for e in elevations:
        for a in range(per):
            cams.append(make_camera(W, H, azim_deg=360.0 * a / per, elev_deg=e))
    return cams[:n] if len(cams) >= n else cams


@dataclass
class Scene:
    """Inputs of the hot path for one avatar (all cages concatenated like models/garment_net.py:149-161)."""

    name: str
    canon_points: np.ndarray  # [V,3] f32
    tetpoints: np.ndarray  # [V,3] f32 deformed cage vertices
    tets: np.ndarray  # [T,4] i64
    tetra_id: np.ndarray  # [N] i64
    barys: np.ndarray  # [N,4] f32
    canonical_gradient: np.ndarray  # [N,3,3] f32  = inv(G(canonical_tetras[tetra_id]))  (lib/cage.py:329)
    scales: np.ndarray  # [N,3] f32 (post-exp)
    rotation: np.ndarray  # [N,4] f32 (r,x,y,z)
    opacity: np.ndarray  # [N,1] f32
    colors: np.ndarray  # [N,3] f32
    shs: np.ndarray | None  # [N,16,3] f32
    W: int = 0
    H: int = 0
    meta: dict = field(default_factory=dict)

    @property
    def N(self):
        return int(self.tetra_id.shape[0])


def _lattice_for(n_tets: int, box):
    """Pick nx,ny,nz ~ proportional to the box with 6*nx*ny*nz ~ n_tets."""
    cubes = max(1, n_tets // 6)
    vol = box[0] * box[1] * box[2]
    h = (vol / cubes) ** (1.0 / 3.0)
    return tuple(max(1, int(round(b / h))) for b in box)


def make_scene(name: str, N: int, n_tets: int, W: int, H: int, seed=0, scale_mul=1.0, garments=(1.0,),
               with_sh=False) -> Scene:
    rng = np.random.default_rng(seed)
    box, origin = (0.6, 1.8, 0.4), (-0.3, -0.9, -0.2)
    pts_l, tets_l, tid_l, voff = [], [], [], 0
    toff = 0
    counts = [int(round(N * g)) for g in garments]
    counts[-1] = N - sum(counts[:-1])
    for gi, (g, n_g) in enumerate(zip(garments, counts)):
        nx, ny, nz = _lattice_for(max(6, int(n_tets * g)), box)
        p, t = make_cage(nx, ny, nz, box, origin)
        if gi > 0:  # garments: slightly inflated copies of the body cage, like layered clothing
            p = (p * (1.0 + 0.04 * gi)).astype(np.float32)
        tid = rng.integers(0, t.shape[0], size=n_g) + toff
        pts_l.append(p)
        tets_l.append(t + voff)
        tid_l.append(tid)
        voff += p.shape[0]
        toff += t.shape[0]
    canon = np.concatenate(pts_l).astype(np.float32)
    tets = np.concatenate(tets_l).astype(np.int64)
    tetra_id = np.concatenate(tid_l).astype(np.int64)
    # Dirichlet(1,1,1,1) barycentrics, strictly inside (min >= 0.02)
    b = rng.dirichlet(np.ones(4), size=N)
    bad = b.min(1) < 0.02
    while bad.any():
        b[bad] = rng.dirichlet(np.ones(4), size=int(bad.sum()))
        bad = b.min(1) < 0.02
    barys = b.astype(np.float32)
    scales = (np.exp(rng.normal(math.log(0.006), 0.35, size=(N, 1))) * rng.uniform(0.5, 1.5, size=(N, 3)) * scale_mul)
    q = rng.normal(size=(N, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opacity = 1.0 / (1.0 + np.exp(-rng.normal(0.0, 1.5, size=(N, 1))))
    colors = rng.uniform(0.0, 1.0, size=(N, 3))
    shs = rng.normal(0.0, 0.3, size=(N, 16, 3)).astype(np.float32) if with_sh else None
    Xc = canon[tets[tetra_id]].astype(np.float64)
    cg = np.linalg.inv(tet_gradient(Xc)).astype(np.float32)
    return Scene(name, canon, deform_points(canon), tets, tetra_id, barys, cg, scales.astype(np.float32),
                 q.astype(np.float32), opacity.astype(np.float32), colors.astype(np.float32), shs, W, H,
                 meta=dict(seed=seed, garments=list(garments), scale_mul=scale_mul))


_GARMENTS = (0.26, 0.37, 0.37)  # body / upper / lower = 35k / 50k / 50k of 135k (configs/actorshq_actor02.yml:123-142)

def sort_scene_by_tet(scene: "Scene") -> "Scene":
    """The same avatar with its Gaussians stored in tet order (a stable argsort of tetra_id applied to every per-Gaussian
    array).  The order of the Gaussians means nothing to the model; stored like this, Gaussians that share a tet -- and
    are therefore close in space -- are neighbours in memory, which the deformation kernels (coalesced instead of
    gathered accesses, in-warp vertex sums) and the blend kernels' gathers reward.  What a user does once after
    `compute_bary` in CageBase.load_tetra (lib/cage.py:308-337)."""
    import dataclasses
    o = np.argsort(scene.tetra_id, kind="stable")
    take = lambda a: None if a is None else np.ascontiguousarray(a[o])  # noqa: E731
    return dataclasses.replace(scene, tetra_id=take(scene.tetra_id), barys=take(scene.barys),
                               canonical_gradient=take(scene.canonical_gradient), scales=take(scene.scales),
                               rotation=take(scene.rotation), opacity=take(scene.opacity), colors=take(scene.colors),
                               shs=take(scene.shs), meta=dict(scene.meta, tet_sorted=True))


WORKLOADS = {
    # name: (N, n_tets, W, H, scale_mul, garments)
    "c1": (100_000, 10_368, 512, 512, 1.0, (1.0,)),
    "c2": (300_000, 30_000, 1024, 1024, 1.0, _GARMENTS),
    "c3": (1_000_000, 100_000, 2048, 1334, 1.0, _GARMENTS),
    "c4": (4_000_000, 400_000, 3840, 2160, 4.0, _GARMENTS),
    # BASELINE.json's metric string: "fwd+bwd views/s @1M Gaussians 1024^2"
    "metric": (1_000_000, 100_000, 1024, 1024, 1.0, _GARMENTS),
    "tiny": (2_000, 384, 128, 96, 2.0, (1.0,)),
}


def make_workload(name: str, seed=0, with_sh=False) -> Scene:
    N, T, W, H, sm, g = WORKLOADS[name]
    return make_scene(name, N, T, W, H, seed=seed, scale_mul=sm, garments=g, with_sh=with_sh)


# ---- synthetic skinned skeleton (cage-vertex LBS) --------------------------------------
def make_skeleton(n_vertices: int = 2197, seed: int = 0, fingers: int = 5, finger_len: int = 4,
                  max_influences: int = 8, box=(600.0, 1800.0, 400.0)):
    """A humanoid-shaped joint tree plus skinned vertices, in the dictionary formats the reference's constructors read.

    Returns (model_json, lbs_config_dict, lbs_scale[1,NS], motion[1,NP]):
      * model_json["Skeleton"]["Bones"][j] = {Name, Parent, PreRotation (x,y,z,w), TranslationOffset};
        model_json["SkinnedModel"] = {RestPositions, RestVertexNormals, SkinningWeights [(joint, w)...],
        SkinningOffsets [V+1], Faces{Indices, TextureIndices}, TextureCoordinates}  (lbsmodel/body_model.py:62-101);
      * lbs_config_dict = {channel_names, transform [7*NJ, NP+NS], transform_offsets [1, 7*NJ], limits,
        nr_scaling_params, nr_position_params}  (lbsmodel/body_model.py:24-38, :562-576).
    Joint 0 is a world joint (Parent beyond the bone count -> -1), joint 1 the body root the reference reads its
    root transform from (body_model.py:190).  Units are millimetres like the reference's body model
    (lib/blueman.py:150 divides by 1000).  ~160 joints with the defaults, 6 global + one parameter per rotating
    channel, scale parameters shared per limb.
    """
    rng = np.random.default_rng(seed)
    names, parents, offs = [], [], []

    def add(name, parent, off):
        names.append(name)
        parents.append(parent)
        offs.append(off)
        return len(names) - 1

    world = add("world", 10 ** 6, (0.0, 0.0, 0.0))
    root = add("b_root", world, (0.0, 900.0, 0.0))
    spine = root
    for i in range(4):
        spine = add(f"b_spine{i}", spine, (0.0, 120.0, 5.0))
    neck = add("b_neck", spine, (0.0, 90.0, 0.0))
    head = add("b_head", neck, (0.0, 110.0, 10.0))
    for e in ("l", "r"):
        add(f"b_{e}_eye", head, (30.0 if e == "l" else -30.0, 40.0, 70.0))
    for side, sx in (("l", 1.0), ("r", -1.0)):
        clav = add(f"b_{side}_clav", spine, (sx * 60.0, 60.0, 0.0))
        arm = add(f"b_{side}_arm", clav, (sx * 130.0, 0.0, 0.0))
        twist = add(f"b_{side}_arm_twist", arm, (sx * 130.0, 0.0, 0.0))
        fore = add(f"b_{side}_forearm", twist, (sx * 130.0, 0.0, 0.0))
        ftw = add(f"b_{side}_forearm_twist", fore, (sx * 120.0, 0.0, 0.0))
        wrist = add(f"b_{side}_wrist", ftw, (sx * 120.0, 0.0, 0.0))
        for f in range(fingers):
            j = wrist
            for k in range(finger_len):
                j = add(f"b_{side}_f{f}_{k}", j, (sx * (40.0 - 5 * k), 0.0, 12.0 * (f - 2) if k == 0 else 0.0))
        up = add(f"b_{side}_upleg", root, (sx * 90.0, -60.0, 0.0))
        uptw = add(f"b_{side}_upleg_twist", up, (0.0, -200.0, 0.0))
        leg = add(f"b_{side}_leg", uptw, (0.0, -220.0, 0.0))
        legtw = add(f"b_{side}_leg_twist", leg, (0.0, -200.0, 0.0))
        foot = add(f"b_{side}_foot", legtw, (0.0, -210.0, -20.0))
        ball = add(f"b_{side}_ball", foot, (0.0, -50.0, 120.0))
        for t in range(5):
            toe = add(f"b_{side}_toe{t}_0", ball, (sx * 12.0 * (t - 2), 0.0, 30.0))
            add(f"b_{side}_toe{t}_1", toe, (0.0, 0.0, 20.0))
    nj = len(names)
    pre = rng.normal(size=(nj, 4)) * np.array([0.15, 0.15, 0.15, 0.0]) + np.array([0.0, 0.0, 0.0, 1.0])
    pre /= np.linalg.norm(pre, axis=1, keepdims=True)
    pre[0] = (0.0, 0.0, 0.0, 1.0)
    offs = np.asarray(offs, np.float64) + rng.normal(size=(nj, 3)) * 2.0
    offs[0] = 0.0
    bones = [{"Name": names[j], "Parent": int(parents[j]), "PreRotation": [float(x) for x in pre[j]],
              "TranslationOffset": [float(x) for x in offs[j]]} for j in range(nj)]

    # rest-pose joint positions (pre-rotations ignored: only used to place vertices and pick influences)
    pos = np.zeros((nj, 3))
    for j in range(1, nj):
        pos[j] = pos[parents[j]] + offs[j] if parents[j] < nj else offs[j]
    lo, hi = pos.min(0) - 60.0, pos.max(0) + 60.0
    verts = rng.uniform(lo, hi, size=(n_vertices, 3))
    weights_flat, offsets = [], [0]
    for v in range(n_vertices):
        d = np.linalg.norm(pos[1:] - verts[v], axis=1)
        k = int(rng.integers(1, max_influences + 1))
        nearest = np.argsort(d, kind="stable")[:k] + 1
        w = np.exp(-d[nearest - 1] / 80.0) + 1e-3
        w /= w.sum()
        weights_flat += [[int(j), float(x)] for j, x in zip(nearest, w)]
        offsets.append(len(weights_flat))
    model_json = {
        "Skeleton": {"Bones": bones},
        "SkinnedModel": {
            "RestPositions": verts.astype(np.float32).tolist(),
            "RestVertexNormals": np.tile(np.array([[0.0, 0.0, 1.0]], np.float32), (n_vertices, 1)).tolist(),
            "SkinningWeights": weights_flat, "SkinningOffsets": offsets,
            "Faces": {"Indices": [0, 1, 2], "TextureIndices": [0, 1, 2]},
            "TextureCoordinates": [0.0, 0.0, 1.0, 0.0, 0.0, 1.0],
        },
    }

    # parameter transform: 6 global parameters drive the root's tx..rz; every other joint gets rx/ry/rz parameters
    # (twist joints only rx); scale parameters are shared by limb and feed the "sc" channel.
    triplets, n_pos = [], 0
    for c in range(6):
        triplets.append((7 * root + c, n_pos, 1.0))
        n_pos += 1
    for j in range(2, nj):
        chans = (3,) if "twist" in names[j] else (3, 4, 5)
        for c in chans:
            triplets.append((7 * j + c, n_pos, 1.0))
            if "twist" in names[j] and n_pos > 6:                # twist joints also follow their parent's parameter
                triplets.append((7 * j + c, n_pos - 1, 0.5))
            n_pos += 1
    groups = ["spine", "neck", "head", "eye", "l_clav", "r_clav", "l_arm", "r_arm", "l_forearm", "r_forearm", "l_wrist",
              "r_wrist", "l_f", "r_f", "l_upleg", "r_upleg", "l_leg", "r_leg", "l_foot", "r_foot", "l_ball", "r_ball",
              "l_toe", "r_toe", "root"]
    n_scale = len(groups)
    for j in range(1, nj):
        for gi, gname in enumerate(groups):
            if gname in names[j]:
                triplets.append((7 * j + 6, n_pos + gi, 1.0))
                break
    transform = np.zeros((7 * nj, n_pos + n_scale), np.float32)
    for r, c, v in triplets:
        transform[r, c] = v
    cfg = {"channel_names": ["tx", "ty", "tz", "rx", "ry", "rz", "sc"], "transform": transform,
           "transform_offsets": np.zeros((1, 7 * nj), np.float32), "limits": [],
           "nr_scaling_params": n_scale, "nr_position_params": n_pos}
    lbs_scale = (rng.normal(size=(1, n_scale)) * 0.05).astype(np.float32)
    motion = (rng.normal(size=(1, n_pos)) * 0.35).astype(np.float32)
    motion[0, :3] = (30.0, -20.0, 45.0)
    return model_json, cfg, lbs_scale, motion
