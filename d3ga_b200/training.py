"""Host-side mirror of one D3GA optimisation step restricted to the hot path (models/trainer.py:91-140,186-192):
cage deformation -> render -> loss -> backward, per camera, gradients accumulated; at the end of a multi-view step
one all-reduce of the flat gradient buffer (d3ga_b200.dist).  Uses only the public drop-in API."""
from __future__ import annotations

import numpy as np
import torch

from d3ga_b200 import dist as ddist
from d3ga_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from d3ga_b200.tetra_sampler import deform_gaussians


class CameraBank:
    """Cameras as one packed pinned host array [n, 35] (view 16 | proj 16 | campos 3) and its device copy."""

    def __init__(self, cams, device):
        self.cams = cams
        packed = np.stack([np.concatenate([c.viewmatrix.ravel(), c.projmatrix.ravel(), c.campos.ravel()]) for c in cams])
        self.host = torch.from_numpy(packed.astype(np.float32))
        if device.type == "cuda":
            self.host = self.host.pin_memory()
        self.dev = self.host.to(device)
        self.slot = torch.empty(35, dtype=torch.float32, device=device)   # e2e: per-view upload target

    def settings(self, i, W, H, bg, from_host=False):
        if from_host:
            self.slot.copy_(self.host[i], non_blocking=True)
            row = self.slot
        else:
            row = self.dev[i]
        return self.settings_from_row(i, row, W, H, bg)

    def settings_from_row(self, i, row, W, H, bg):
        """Settings of camera i whose 35 floats (view | proj | campos) are in the device tensor `row`."""
        c = self.cams[i]
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=c.tanfovx, tanfovy=c.tanfovy, bg=bg, scale_modifier=1.0,
            viewmatrix=row[0:16].view(4, 4), projmatrix=row[16:32].view(4, 4), sh_degree=0, campos=row[32:35],
            prefiltered=False, debug=False, antialiasing=False)


class Avatar:
    """Resident parameters of the hot path (what nn.Parameters / buffers hold in models/cage_net.py)."""

    def __init__(self, scene, device):
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)  # noqa: E731
        self.device = device
        self.W, self.H, self.N = scene.W, scene.H, scene.N
        self.tets, self.tetra_id, self.cg = t(scene.tets), t(scene.tetra_id), t(scene.canonical_gradient)
        self.tetpoints = t(scene.tetpoints).requires_grad_(True)
        self.barys = t(scene.barys).requires_grad_(True)
        # the scale PARAMETER is pre-activation like the reference's (models/cage_net.py:66-70: log of a length; the exp of
        # cage_net.py:214 runs inside the fused deformation launch)
        self.scaling = t(np.log(scene.scales)).requires_grad_(True)
        self.rotation = t(scene.rotation).requires_grad_(True)
        self.opacity = t(scene.opacity).requires_grad_(True)
        self.colors = t(scene.colors).requires_grad_(True)
        self.silhouette = torch.ones_like(self.colors)   # models/cage_net.py:249 silhouette_rgb
        self.params = [self.tetpoints, self.barys, self.scaling, self.rotation, self.opacity, self.colors]
        # inverse rest-pose edge matrices of the cage energy (lib/cage.py:312-313)
        canon = t(scene.canon_points)
        X = canon[self.tets]
        self.dn_inv = torch.linalg.inv(torch.stack([X[:, 3] - X[:, 0], X[:, 2] - X[:, 0], X[:, 1] - X[:, 0]], dim=2)).contiguous()
        self.grads = ddist.FlatGrads(self.params)
        self._means2D = None
        self.tetpoints_host = torch.from_numpy(scene.tetpoints.copy())
        if device.type == "cuda":
            self.tetpoints_host = self.tetpoints_host.pin_memory()

    def shadow(self):
        """A second handle on the same parameters with its OWN gradient buffer and its own copy of the per-view input
        `tetpoints`: what a view needs to run on another stream concurrently with a view of `self`."""
        import copy
        o = copy.copy(self)
        o.tetpoints = self.tetpoints.detach().clone().requires_grad_(True)
        o.barys, o.scaling, o.rotation, o.opacity, o.colors = (
            p.detach().requires_grad_(True) for p in (self.barys, self.scaling, self.rotation, self.opacity, self.colors))
        o.params = [o.tetpoints, o.barys, o.scaling, o.rotation, o.opacity, o.colors]
        o.grads = ddist.FlatGrads(o.params)
        return o

    def render(self, rs):
        """What GarmentNet.forward + render() do on the hot path (cage_net.py:213-230, renderer.py:122-141)."""
        means3D, cov6 = deform_gaussians(self.tetpoints, self.tets, self.tetra_id, self.barys, self.cg, self.scaling,
                                         self.rotation, log_scales=True)
        # renderer.py:122-126 allocates zeroed screen-space points per call only to receive a gradient nobody reads on
        # this path; a resident zero buffer serves the API without two extra kernels per view
        if self._means2D is None or self._means2D.shape != means3D.shape:
            self._means2D = torch.zeros_like(means3D)
        means2D = self._means2D
        return GaussianRasterizer(raster_settings=rs)(
            means3D=means3D, means2D=means2D, shs=None, colors_precomp=self.colors, opacities=self.opacity,
            scales=None, rotations=None, cov3D_precomp=cov6)[0]


def view_step(avatar, rs, weights):
    """One view forward + backward; loss = sum(image * weights).

    dL/dimage of this loss IS `weights`, so the backward starts from the image with `weights` as the incoming gradient:
    no ones-fill and no weights * 1 product ([3,H,W] each) per view; the loss value is one dot product beside it.
    (Measured on one box, 4 views in flight: 1469 -> 1492 views/s.  Adding the six parameter gradients into the flat
    buffer with one multi-tensor kernel instead of autograd's six in-place adds was measured too: 1374, worse.)"""
    img = avatar.render(rs)
    loss = torch.dot(img.detach().reshape(-1), weights.reshape(-1))
    img.backward(weights)
    return loss


def frame_pair_step(avatar, rs, weights, fem_weight=1.0):
    """One D3GA training FRAME with the settings `rs` (models/trainer.py:91-140): the deformation block with the cage's
    energy out of the same launch, the RGB render over rs.bg and the silhouette render over black as ONE
    GaussianRasterizer.forward_pair call, and one backward through both images and the energy term.  Same signature as
    view_step, so GraphedViewStep / PipelinedViews can capture and replay it (step=frame_pair_step)."""
    means3D, cov6, energy = deform_gaussians(avatar.tetpoints, avatar.tets, avatar.tetra_id, avatar.barys, avatar.cg,
                                             avatar.scaling, avatar.rotation, log_scales=True, dn_inv=avatar.dn_inv)
    if avatar._means2D is None or avatar._means2D.shape != means3D.shape:
        avatar._means2D = torch.zeros_like(means3D)
    img, sil, _ = GaussianRasterizer(raster_settings=rs).forward_pair(
        means3D=means3D, means2D=avatar._means2D, opacities=avatar.opacity, colors_b=avatar.silhouette,
        bg_b=torch.zeros_like(rs.bg), colors_precomp=avatar.colors, cov3D_precomp=cov6)
    loss = (torch.dot(img.detach().reshape(-1), weights.reshape(-1)) + torch.dot(sil.detach().reshape(-1), weights.reshape(-1))
            + fem_weight * energy.detach().mean())
    torch.autograd.backward([img, sil, energy], [weights, weights, torch.full_like(energy, fem_weight / energy.numel())])
    return loss


U8_SCALE, U8_OFFSET = 2.0 / 255.0, -1.0   # a uint8 target / weight image stands for u8 * U8_SCALE + U8_OFFSET
_u8_offset: dict = {}


def dequantise_into(dst, img):
    """dst[3,H,W] f32 <- img: float32 is copied; uint8 is normalised on the device in one kernel, which is what the
    reference does with the camera images it uploads (lib/batch.py:86-88: uint8 over PCIe, float on the GPU)."""
    if img.dtype == torch.uint8:
        off = _u8_offset.get(dst.device)
        if off is None:
            off = _u8_offset[dst.device] = torch.full((), U8_OFFSET, dtype=torch.float32, device=dst.device)
        torch.add(off, img, alpha=U8_SCALE, out=dst)
    elif img.data_ptr() != dst.data_ptr():
        dst.copy_(img, non_blocking=True)
    return dst


class HostFeed:
    """Per-view inputs that arrive from pinned host memory (the e2e arm of bench.py): the target / weight image (float32,
    or uint8 like the reference's camera frames: a quarter of the bytes), the camera (35 floats) and the cage vertices
    are copied on a side stream ONE VIEW AHEAD into a double buffer, so the uploads (3.1 MB as uint8 at 1024^2) overlap the
    previous view's kernels and no host-to-device DMA sits on the compute stream -- what a data loader with a prefetch
    queue does (train.py:426-434 uses pinned DataLoader memory)."""

    def __init__(self, weights_host, device, cam_host=None, tetpoints_host=None, slots=2):
        self.host = weights_host
        self.cam_host, self.tet_host = cam_host, tetpoints_host
        self.stream = torch.cuda.Stream(device=device)
        self.slots = slots       # views in flight + views being uploaded (2 lanes of PipelinedViews want 4)
        self.bufs = [torch.empty_like(weights_host, device=device) for _ in range(slots)]
        self.cams = [torch.empty(35, dtype=torch.float32, device=device) for _ in range(slots)]
        self.tets = [None if tetpoints_host is None else torch.empty_like(tetpoints_host, device=device)
                     for _ in range(slots)]
        self.ready = [torch.cuda.Event() for _ in range(slots)]
        self.free = [torch.cuda.Event() for _ in range(slots)]
        for e in self.free:
            e.record()
        self.k = 0

    def prefetch(self, cam_id=None):
        """Enqueue the upload of the next view's inputs; returns the slot they will land in."""
        slot = self.k % self.slots
        self.k += 1
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.free[slot])      # the view that used this buffer has finished with it
            self.bufs[slot].copy_(self.host, non_blocking=True)
            if cam_id is not None and self.cam_host is not None:
                self.cams[slot].copy_(self.cam_host[cam_id], non_blocking=True)
            if self.tet_host is not None:
                self.tets[slot].copy_(self.tet_host, non_blocking=True)
            self.ready[slot].record(self.stream)
        return slot

    def acquire(self, slot):
        torch.cuda.current_stream().wait_event(self.ready[slot])
        return self.bufs[slot]

    def release(self, slot):
        self.free[slot].record()


def multi_view_step(avatar, bank, cam_ids, weights, bg, from_host=False, feed=None, loss_host=None):
    """`cam_ids` views on this rank, then one all-reduce of the accumulated gradients.  With from_host, every view's
    camera, cage vertices and target image come from pinned host memory and the step's loss is read back at the end."""
    avatar.grads.zero_()
    total = None
    slot = feed.prefetch(cam_ids[0]) if from_host else None
    for n, i in enumerate(cam_ids):
        if from_host:
            weights = feed.acquire(slot)
            if weights.dtype == torch.uint8:
                weights = dequantise_into(torch.empty(weights.shape, dtype=torch.float32, device=weights.device), weights)
            with torch.no_grad():
                if feed.tet_host is not None:
                    avatar.tetpoints.copy_(feed.tets[slot], non_blocking=True)
                else:
                    avatar.tetpoints.copy_(avatar.tetpoints_host, non_blocking=True)
            nxt = feed.prefetch(cam_ids[n + 1]) if n + 1 < len(cam_ids) else None
        if from_host and feed.cam_host is not None:
            rs = bank.settings_from_row(i, feed.cams[slot], avatar.W, avatar.H, bg)
        else:
            rs = bank.settings(i, avatar.W, avatar.H, bg, from_host=from_host)
        loss = view_step(avatar, rs, weights)
        if from_host:
            feed.release(slot)
            slot = nxt
        total = loss.detach() if total is None else total + loss.detach()
    avatar.grads.all_reduce()
    if from_host:
        loss_host.copy_(total, non_blocking=True)     # the step's result goes back to the host
        torch.cuda.current_stream().synchronize()
    return total


class GraphedViewStep:
    """`view_step` captured once into a CUDA graph and replayed per camera.

    The eager per-view path costs about a millisecond of Python, autograd and launch overhead -- as much as the kernels
    themselves at 1 M Gaussians and more than them below that -- and waits for the binning count in the middle of every
    view.  The captured graph holds the same launches (deformation, the rasterizer's two stages, the loss, the whole
    backward) with static inputs: the camera is a 35-float device buffer (view | proj | campos) the rasterization
    settings alias, the target / weight image a fixed [3,H,W] buffer, the parameters and their flat gradient buffer the
    avatar's own tensors.  Intrinsics (tanfov) and the background are baked in at capture.

    The binning workspace is sized from the capacity hint at capture time; every replay folds its (num_rendered,
    overflow) into a device counter and `finish()` reads it back: if a view overflowed, the hint has been raised, the
    graph must be captured again (`capture()`) and the step redone -- `multi_view_step(..., graphed=...)` does that."""

    def __init__(self, avatar, cam, bg, weights=None, views_in_flight=1, step=None):
        from d3ga_b200 import diff_gaussian_rasterization as dgr
        self.dgr = dgr
        self.avatar = avatar
        self.views_in_flight = int(views_in_flight)   # how many such graphs the caller replays concurrently
        self.step = view_step if step is None else step   # what one replay does: (avatar, settings, weights) -> loss
        dev = avatar.device
        self.cam = torch.zeros(35, dtype=torch.float32, device=dev)
        self.weights = torch.empty((3, avatar.H, avatar.W), dtype=torch.float32, device=dev) if weights is None else weights
        self.total = torch.zeros((), dtype=torch.float32, device=dev)     # sum of the losses since reset()
        self.rs = GaussianRasterizationSettings(
            image_height=avatar.H, image_width=avatar.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=bg,
            scale_modifier=1.0, viewmatrix=self.cam[0:16].view(4, 4), projmatrix=self.cam[16:32].view(4, 4), sh_degree=0,
            campos=self.cam[32:35], prefiltered=False, debug=False, antialiasing=False)
        self.graph = None
        self.capture_stream = torch.cuda.Stream(device=dev)   # not torch's shared default: the deferred overflow
        self.launches_per_replay = 0                          # counter is kept per capture stream
        self.replays = 0
        self.captures = 0

    def capture(self, cam_row, warm=True):
        """Warm up eagerly on a side stream (sizes the capacity hint, fills the allocator), then capture one view.
        warm=False captures against the capacity hint as it stands (tests use it to provoke an overflow)."""
        from d3ga_b200 import _lib
        self.cam.copy_(cam_row, non_blocking=True)
        if warm:
            # an eager view sets the capacity hint from ITS duplicate count; the graph must be sized for the largest view
            # seen so far (collect_deferred raised the hint after an overflow), so the warm-up may only raise it
            hkey = (self.avatar.device, self.avatar.N, self.avatar.W, self.avatar.H)
            hint = self.dgr._capacity_hint.get(hkey, 0)
            keep = self.avatar.grads.flat.clone()        # the warm-up views must not leak into the caller's gradients
            side = torch.cuda.Stream(device=self.avatar.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    self.step(self.avatar, self.rs, self.weights)
            torch.cuda.current_stream().wait_stream(side)
            self.avatar.grads.flat.copy_(keep)
            self.dgr._capacity_hint[hkey] = max(hint, self.dgr._capacity_hint.get(hkey, 0))
        torch.cuda.synchronize()
        # the overflow counter of this lane must exist (and be zero) before the capture: created inside it, its zero fill
        # would be replayed with every view and only the last view of a step would ever be seen
        self.dgr.prepare_deferred(self.avatar.device, self.capture_stream)
        self.graph = torch.cuda.CUDAGraph()
        l0 = _lib.launch_count()
        # the launch configurations baked into the graph are sized for `views_in_flight` concurrent replays (the direct
        # tile scatter then takes about 1/n of the GPU for n times as long); eager calls keep the single-view sizing
        lib = _lib.load()
        lib.d3ga_set_views_in_flight(self.views_in_flight)
        try:
            with self.dgr.deferred_overflow_check(), torch.cuda.graph(self.graph, stream=self.capture_stream):
                loss = self.step(self.avatar, self.rs, self.weights)
                self.total += loss.detach()
        finally:
            lib.d3ga_set_views_in_flight(1)
        self.launches_per_replay = _lib.launch_count() - l0
        self.captures += 1

    def run(self, cam_row, weights=None):
        """One view: cam_row [35] (device or pinned host), weights None = already in self.weights."""
        from d3ga_b200 import _lib
        if self.graph is None:
            self.capture(cam_row)
        self.cam.copy_(cam_row, non_blocking=True)
        if weights is not None:
            dequantise_into(self.weights, weights)
        self.graph.replay()
        self.replays += 1
        _lib.note_graph_launches(self.launches_per_replay)

    def reset(self):
        self.total.zero_()

    def finish(self):
        """Waits for the replays issued so far; returns (largest num_rendered, overflowed)."""
        return self.dgr.collect_deferred(self.avatar.device)


class PipelinedViews:
    """`lanes` GraphedViewSteps on as many streams, used round-robin: consecutive views of a step are independent given
    the parameters, so view n+1's preprocess / sort / emission fill the SMs that view n's blend kernels leave idle in
    their tails (the heaviest tiles' serial chains) and the latency-bound binning kernels overlap each other.  Each lane
    works on a shadow of the avatar (same parameter storage, own gradient buffer and own cage-vertex input); `finish()`
    adds the lanes' gradients into the avatar's flat buffer."""

    @staticmethod
    def lanes_that_fit(avatar, wanted, fraction=0.5):
        """Lanes whose workspaces fit in `fraction` of the free device memory.  A lane keeps a geometry + binning + image
        workspace, the deformation outputs and the backward temporaries alive in its graph pool, and the eager warm-up
        of a capture needs the same again: about 2.5 x the workspace bytes at the current capacity hint (an eager view
        must have run, so that the hint reflects the workload: the 4 M-Gaussian 4K stress needs ~40 GB per lane)."""
        from d3ga_b200 import _lib
        from d3ga_b200 import diff_gaussian_rasterization as dgr
        cap = dgr._capacity_hint.get((avatar.device, avatar.N, avatar.W, avatar.H), 6 * avatar.N)
        per_lane = 2.5 * sum(_lib.workspace_bytes(avatar.N, avatar.W, avatar.H, cap)) + 400 * avatar.N
        free, _ = torch.cuda.mem_get_info(avatar.device)
        return int(max(1, min(wanted, fraction * free // per_lane)))

    def __init__(self, avatar, cam, bg, lanes=2, step=None):
        self.avatar = avatar
        self.shadows = [avatar.shadow() for _ in range(lanes)]
        self.streams = [torch.cuda.Stream(device=avatar.device) for _ in range(lanes)]
        self.gvs = [GraphedViewStep(sh, cam, bg, views_in_flight=lanes, step=step) for sh in self.shadows]

    @property
    def captures(self):
        return sum(g.captures for g in self.gvs)

    def capture(self, cam_row):
        for g, st in zip(self.gvs, self.streams):
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                g.capture(cam_row)
            torch.cuda.current_stream().wait_stream(st)

    def begin(self, cam_row):
        if self.gvs[0].graph is None:
            self.capture(cam_row)
        main = torch.cuda.current_stream()
        for g, sh, st in zip(self.gvs, self.shadows, self.streams):
            st.wait_stream(main)
            with torch.cuda.stream(st):
                sh.grads.zero_()
                g.reset()

    def lane(self, n):
        k = n % len(self.gvs)
        return self.gvs[k], self.shadows[k], self.streams[k]

    def finish(self):
        """Joins the lanes; returns (sum of losses, overflowed)."""
        main = torch.cuda.current_stream()
        for st in self.streams:
            main.wait_stream(st)
        # lanes' gradients -> the avatar's flat buffer -> sum over ranks, chunk by chunk (dist.FlatGrads)
        self.avatar.grads.sum_lanes_and_ranks([sh.grads.flat for sh in self.shadows])
        total = self.gvs[0].total.clone()
        for g in self.gvs[1:]:
            total = total + g.total
        _, overflowed = self.gvs[0].finish()
        return total, overflowed


def multi_view_step_pipelined(avatar, bank, cam_ids, pv, weights=None, from_host=False, feed=None, loss_host=None):
    """multi_view_step over a PipelinedViews: view n runs on lane n % lanes.  Same result as the eager step up to the
    order of the floating-point gradient sums.  With from_host the feed needs 2 * lanes slots: `lanes` views in flight
    plus `lanes` being uploaded."""
    lanes = len(pv.gvs)
    while True:
        pv.begin(bank.dev[cam_ids[0]])
        queue = []
        if from_host:
            assert feed.slots >= 2 * lanes, "HostFeed(slots=2*lanes) for pipelined views"
            queue = [feed.prefetch(i) for i in cam_ids[:lanes]]
        for n, i in enumerate(cam_ids):
            gv, sh, st = pv.lane(n)
            with torch.cuda.stream(st):
                w, row = weights, bank.dev[i]
                if from_host:
                    slot = queue.pop(0)
                    w = feed.acquire(slot)
                    with torch.no_grad():
                        sh.tetpoints.copy_(feed.tets[slot] if feed.tet_host is not None else avatar.tetpoints_host,
                                           non_blocking=True)
                    row = feed.cams[slot] if feed.cam_host is not None else bank.host[i]
                    if n + lanes < len(cam_ids):
                        queue.append(feed.prefetch(cam_ids[n + lanes]))
                gv.run(row, w)
                if from_host:
                    feed.release(slot)
        total, overflowed = pv.finish()      # includes the all-reduce over ranks
        if from_host:
            loss_host.copy_(total, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        if not overflowed:
            return total
        pv.capture(bank.dev[cam_ids[0]])     # the capacity hint now covers the largest view seen


def multi_view_step_graphed(avatar, bank, cam_ids, gv, weights=None, from_host=False, feed=None, loss_host=None):
    """multi_view_step through a GraphedViewStep: same work per view, one host synchronisation per step (the overflow
    check; with from_host also the loss read-back).  Redoes the step after re-capturing if a view overflowed."""
    while True:
        avatar.grads.zero_()
        gv.reset()
        slot = feed.prefetch(cam_ids[0]) if from_host else None
        for n, i in enumerate(cam_ids):
            w, row = weights, bank.dev[i]
            if from_host:
                w = feed.acquire(slot)
                with torch.no_grad():
                    if feed.tet_host is not None:
                        avatar.tetpoints.copy_(feed.tets[slot], non_blocking=True)
                    else:
                        avatar.tetpoints.copy_(avatar.tetpoints_host, non_blocking=True)
                row = feed.cams[slot] if feed.cam_host is not None else bank.host[i]
                nxt = feed.prefetch(cam_ids[n + 1]) if n + 1 < len(cam_ids) else None
            gv.run(row, w)
            if from_host:
                feed.release(slot)
                slot = nxt
        avatar.grads.all_reduce()
        if from_host:
            loss_host.copy_(gv.total, non_blocking=True)
        _, overflowed = gv.finish()
        if not overflowed:
            return gv.total
        gv.capture(bank.dev[cam_ids[0]])     # the capacity hint now covers the largest view seen


def frame_step(avatar, bank, cam_id, weights, bg, fem_weight=1.0, fused_pair=False):
    """One D3GA training frame on the hot path (models/trainer.py:91-140): the deformation block once -- with the cage's
    neo-Hookean energy of cage_net.py:225 out of the same launch --, then TWO renders of the same Gaussians -- RGB on a
    random background and the silhouette (constant colour, black background), each through its own render() call with its
    own Camera object -- and one backward through both plus the energy term (train.py:186-244 adds it to the loss).
    `fused_pair`: the two renders as ONE GaussianRasterizer.forward_pair call (one binning, one blend each way)."""
    means3D, cov6, energy = deform_gaussians(avatar.tetpoints, avatar.tets, avatar.tetra_id, avatar.barys, avatar.cg,
                                             avatar.scaling, avatar.rotation, log_scales=True, dn_inv=avatar.dn_inv)
    if fused_pair:
        rs = bank.settings(cam_id, avatar.W, avatar.H, bg)
        means2D = torch.zeros_like(means3D, requires_grad=True) + 0
        img, sil, _ = GaussianRasterizer(raster_settings=rs).forward_pair(
            means3D=means3D, means2D=means2D, opacities=avatar.opacity, colors_b=avatar.silhouette,
            bg_b=torch.zeros_like(bg), colors_precomp=avatar.colors, cov3D_precomp=cov6)
        loss = (img * weights).sum() + (sil * weights).sum() + fem_weight * energy.mean()
        loss.backward()
        return loss
    imgs = []
    for colors, background in ((avatar.colors, bg), (avatar.silhouette, torch.zeros_like(bg))):
        rs = bank.settings(cam_id, avatar.W, avatar.H, background)
        rs = rs._replace(viewmatrix=rs.viewmatrix.clone(), projmatrix=rs.projmatrix.clone(), campos=rs.campos.clone())
        means2D = torch.zeros_like(means3D, requires_grad=True) + 0
        imgs.append(GaussianRasterizer(raster_settings=rs)(
            means3D=means3D, means2D=means2D, shs=None, colors_precomp=colors, opacities=avatar.opacity, scales=None,
            rotations=None, cov3D_precomp=cov6)[0])
    loss = (imgs[0] * weights).sum() + (imgs[1] * weights).sum() + fem_weight * energy.mean()
    loss.backward()
    return loss
