"""Feature-stage training step on the B200 op (the "train iter/s" half of the metric).

Mirrors ``train.py --include_feature`` of the reference, one iteration per ``step()``:

    render (include_feature)          /root/reference/train.py:147, gaussian_renderer/__init__.py:95-98,108-119
      logits -> softmax/top-k/renorm  scene/gaussian_model.py:510-518, utils/vq_utils.py:9-24
    decode one codebook layer         scene/gaussian_model.py:533-543 (layer_num = 1 in the paper, train.py:154)
    normalise, cosine (+ L1) loss     train.py:157-167, utils/loss_utils.py:18-25
    backward                          train.py:173
    Adam on (logits, codebooks)       scene/gaussian_model.py:235-236,255 ; train.py:261-263

Geometry is frozen (scene/gaussian_model.py:238-243).  Data parallel: each rank renders its own view and the
gradients of (logits, codebooks) are summed over the ranks over one flat buffer, in 4 row blocks pipelined with the
top-k backward: reduce-scatter -> Adam on this rank's shard -> all-gather of the parameters (``dist.ShardedExchange``),
SUM as in the reference's ``--accum_iter`` (train.py:261-263,348).  Every kernel of the step (top-k fwd/bwd,
preprocess, binning, blend fwd/bwd, Gram loss, codebook GEMMs, Adam) is liblsv2's.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from . import dist as lsdist
from .decode import decode_feature_map
from .loss import column_sumsq, decode_cos_loss, decode_cos_loss_segments, normalized_cos_loss
from .optim import FlatAdam
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer
from . import sparse as lssparse
from .sparse import softmax_to_topk_soft_code


def cos_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """utils/loss_utils.py:24-25."""
    return 1 - F.cosine_similarity(network_output, gt, dim=0).mean()


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """utils/loss_utils.py:18-19."""
    return torch.abs(network_output - gt).mean()


class FeatureTrainer:
    """Resident scene + the two trainable tensors + Adam; ``step(cam, gt, mask)`` is one train.py iteration.

    Parameters, gradients and Adam moments live in four flat fp32 buffers ``[logits | codebooks | pad]``
    (``self.logits`` / ``self.codebooks`` are views), so the optimizer is ``optim.FlatAdam`` (one kernel per range) and
    the data-parallel exchange works on flat ranges.  ``collective``:
      "zero1"      reduce-scatter -> each rank steps its 1/world shard -> all-gather of the parameters (default, N > 1)
      "allreduce"  one SUM all-reduce of the gradient, every rank steps everything (the round-1 scheme)
    both in row blocks pipelined with the top-k backward.  ``use_graph=True`` (single rank) captures the steady-state
    iteration of each (view_id, layer) in a CUDA graph: ``step(..., view_id=i)`` replays it."""

    def __init__(self, scene: Dict, topk: int = 4, lr: float = 0.0025, device="cuda", normalize: bool = True,
                 use_cos: bool = True, use_l1: bool = False, layers: Optional[int] = None,
                 fused_loss: bool = True, gram_loss: bool = True, collective: str = "zero1",
                 use_graph: bool = False, check_every: int = 64, defer_codebook_grad: Optional[bool] = None,
                 geometry_cache_gb: float = 0.0, fused_rows: Optional[bool] = None):
        dev = torch.device(device)
        f32 = lambda t: t.to(dev, torch.float32).contiguous()
        self.dev = dev
        self.topk = topk
        # one level of 64 codes: the top-k kernel writes the compacted + packed rows too (LSV2_FUSED_ROWS=0: A/B switch)
        self.fused_rows = (os.environ.get("LSV2_FUSED_ROWS", "1") != "0") if fused_rows is None else bool(fused_rows)
        self.normalize, self.use_cos, self.use_l1 = normalize, use_cos, use_l1
        self.fused_loss = fused_loss        # cos loss (+ the normalisation) as two kernels instead of torch's ~28 passes
        self.gram_loss = gram_loss          # ... and, for a single layer, without forming the [512, H, W] feature map
        self._g2_by_view = {}               # |gt_p|^2 per caller-supplied view id (never keyed by address)
        self.xyz, self.opacity = f32(scene["xyz"]), f32(scene["opacity"])
        self.scales, self.rotations = f32(scene["scaling"]), f32(scene["rotation"])
        self.shs = f32(scene["features"])
        self.sh_degree = int(scene.get("sh_degree", 3))
        cb = f32(scene["codebooks"])
        self.layers = int(cb.shape[0] if layers is None else layers)
        self.codebook_size = int(cb.shape[1])
        lg = f32(scene["logits"])[:, : self.layers * self.codebook_size]
        cb = cb[: self.layers]
        self.rank, self.world, _ = lsdist.world_info()
        if self.world > 1 and not (torch.distributed.is_available() and torch.distributed.is_initialized()):
            self.world, self.rank = 1, 0
        if collective not in ("zero1", "allreduce"):
            raise ValueError("collective must be 'zero1' or 'allreduce'")
        # ---- flat buffers: [logits | codebooks | pad to a multiple of 4 * world] ----
        n_log, n_cb = lg.numel(), cb.numel()
        unit = 4 * self.world
        tail = -(-n_cb // unit) * unit
        self.n_log, self.n_cb = n_log, n_cb
        self.flat_p = torch.zeros(n_log + tail, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros_like(self.flat_p)
        self.flat_p[:n_log].copy_(lg.reshape(-1))
        self.flat_p[n_log:n_log + n_cb].copy_(cb.reshape(-1))
        self.logits = torch.nn.Parameter(self.flat_p[:n_log].view(lg.shape))
        self.codebooks = torch.nn.Parameter(self.flat_p[n_log:n_log + n_cb].view(cb.shape))
        self.logits.grad = self.flat_g[:n_log].view(lg.shape)
        self.codebooks.grad = self.flat_g[n_log:n_log + n_cb].view(cb.shape)
        # the reference: torch.optim.Adam(l, lr=0.0, eps=1e-15) with the group lr = language_feature_lr
        self.optimizer = FlatAdam(self.flat_p, self.flat_g, lr=lr, eps=1e-15) if dev.type == "cuda" else None
        self.bg = torch.zeros(3, dtype=torch.float32, device=dev)
        self.means2D = torch.zeros_like(self.xyz)                      # grad sink of the reference API; unused here
        # ---- row blocks of the logits (+ one block for the codebooks): block i's gradient leaves the top-k backward
        # and its exchange runs under block i + 1's kernel ----
        self.overlap_chunks = int(os.environ.get("LSV2_OVERLAP_CHUNKS", "4")) if self.world > 1 else 1
        C = lg.shape[1]
        n = lg.shape[0]
        rows = -(-(-(-n // max(self.overlap_chunks, 1))) // 1024) * 1024
        self._chunk_rows = [(a, min(n, a + rows)) for a in range(0, n, rows)]
        self._blocks = [(a * C, b * C) for a, b in self._chunk_rows] + [(n_log, n_log + tail)]
        # the sink needs whole-row blocks written by ONE top-k backward over all the logits columns
        self._sink_ok = self.layers == 1 and dev.type == "cuda"
        self.collective = "none"
        self.exchange = None
        self.bucket = None
        if self.world > 1:
            ok = all((b - a) % unit == 0 and a % 4 == 0 for a, b in self._blocks) and self._sink_ok
            self.collective = collective if (collective == "allreduce" or ok) else "allreduce"
            if self.collective == "zero1":
                self.exchange = lsdist.ShardedExchange(self.flat_g, self.flat_p, self._blocks)
            else:
                self.bucket = lsdist.GradBucket([self.logits, self.codebooks])   # only its slice machinery is used
                self.bucket.flat = self.flat_g
                self.bucket.attached = True
        # ---- coefficient gradient in row form (one level of 64 codes; LSV2_ROW_GRAD=0: A/B switch).  The blend backward
        # writes [N, k] instead of a dense [N, 64] array; the top-k backward is linear in it, so at N > 1 the ranks
        # reduce-scatter the ROWS (16 B per Gaussian instead of 256 B) and each runs the top-k backward + Adam on its
        # own shard of rows only ----
        self.row_grad = (self.fused_rows and self._sink_ok and lg.shape[1] == 64 and self.topk <= 16 and
                         os.environ.get("LSV2_ROW_GRAD", "1") != "0" and self.collective in ("none", "zero1"))
        self._row_stride = -(-self.topk // 4) * 4
        self._rows_g = None
        self.rows_exchange = None
        self._row_ctx = None
        if self.row_grad:
            self._rows_g = torch.zeros((n, self._row_stride), dtype=torch.float32, device=dev)
            if self.collective == "zero1":
                try:
                    self.rows_exchange = lsdist.ShardedExchange(
                        self._rows_g.view(-1), None, [(a * self._row_stride, b * self._row_stride)
                                                      for a, b in self._chunk_rows])
                except ValueError:
                    self.row_grad = False
        # the codebook gradient (two pixel-axis GEMMs nobody waits for until the optimizer) can be computed on a side
        # stream after the rest of the backward has been enqueued (loss.CodebookGradSink): at N > 1 the logits' gradient
        # exchange runs under it; at N = 1 it overlaps the tail of the blend backward and the top-k backward
        # (5.09 -> 4.94 ms per iteration at 1M Gaussians / 1008x756)
        if defer_codebook_grad is None:
            env = os.environ.get("LSV2_DEFER_CB")                  # measurement switch, read once at construction
            defer_codebook_grad = (env != "0")
        self._cb_sink = None
        if defer_codebook_grad and dev.type == "cuda" and self.gram_loss and self.layers == 1:
            from .loss import CodebookGradSink
            self._cb_sink = CodebookGradSink(self.codebooks.grad[0:1])
        # N > 1, zero1: the all-gather of the updated parameters is deferred to the START of the next iteration, where it
        # runs under that iteration's preprocess + binning (which read no trainable parameter: the geometry is frozen).
        # Between iterations a rank's copy of the shards it does not own is one step old: call sync_params() before
        # reading logits / codebooks from outside (checkpoints, evaluation).
        self.defer_gather = (self.collective == "zero1" and os.environ.get("LSV2_DEFER_AG", "1") != "0")
        self.split_prepare = dev.type == "cuda" and os.environ.get("LSV2_SPLIT_PREPARE", "1") != "0"
        # Geometry-stage cache (opt-in).  The feature stage freezes the geometry (scene/gaussian_model.py:238-243), so a
        # view's preprocess + binning result is the same in every iteration that renders it: what the blend and its
        # backward read of it (radii, the packed blend record, the sorted id list, the tile ranges: ~90 MB per view at 1M
        # Gaussians / 1008x756) can stay resident -- a 300-view scene is 27 GB of a B200's 180 GB -- and the iteration
        # then starts at the top-k kernel.  cache_geometry(cam, view_id) fills it; step(view_id=...) uses it.
        self.geometry_cache_bytes = int(geometry_cache_gb * 1e9)
        self._geom_cache = {}
        self._geom_cache_used = 0
        self.geometry_cache_hits = 0
        self._tiny_scratch = torch.zeros(4096, dtype=torch.uint8, device=dev) if dev.type == "cuda" else None
        self.keep_last_grad = False
        self.last_grad = None
        self.iteration = 0
        # instance capacity: the first `calib_steps` iterations read num_rendered back (the reference's one host sync
        # per frame); after that the frame runs with a bounded instance buffer and no host sync.  Overflow is sticky on
        # the device (every frame ORs into one word) and read every `check_every` iterations: the capacity then grows.
        self.capacity: Optional[int] = None
        self.calib_steps = 3
        self.check_every = int(check_every)
        self.headroom = 1.5
        self._max_seen = 0
        self._last_state = None
        self._sticky = torch.zeros(2, dtype=torch.int32, device=dev)   # [overflow seen, max instances seen]
        self.overflow_events = 0
        # at N > 1 the captured iteration contains the NCCL reduce-scatter / all-gather (side stream, joined back)
        self.use_graph = bool(use_graph) and dev.type == "cuda"
        self._graphs = {}
        self._graph_pool = None
        self._eager_iters = 0

    # ---- reference: GaussianModel.get_render_weights (scene/gaussian_model.py:510-518) ----
    def render_weights(self) -> torch.Tensor:
        cs = self.codebook_size
        w = [softmax_to_topk_soft_code(self.logits[:, i * cs:(i + 1) * cs], self.topk) for i in range(self.layers)]
        return w[0] if len(w) == 1 else torch.cat(w, dim=-1)

    def _geometry_stage(self, rs):
        from .rasterizer import rasterize_forward
        with torch.no_grad():
            return rasterize_forward(rs._replace(include_feature=False), self.xyz, self.opacity, shs=self.shs,
                                     scales=self.scales, rotations=self.rotations, capacity=self.capacity, render=False)

    def cache_geometry(self, cam: Dict, view_id) -> bool:
        """Run the geometry stage of ``cam`` once and keep what the blend needs of it under ``view_id`` (see
        ``geometry_cache_gb``).  Returns False when the budget is exhausted or the capacity is not calibrated yet.
        Not to be called under stream capture."""
        from .rasterizer import FrameState
        if self.geometry_cache_bytes <= 0 or self.capacity is None or view_id is None:
            return False
        if view_id in self._geom_cache:
            return True
        full = self._geometry_stage(self.settings(cam))
        keep = FrameState(P=full.P, H=full.H, W=full.W, mode=full.mode, L=full.L, K=full.K, view=full.view,
                          status=full.status, capacity=full.capacity, keep=full.keep, scene=full.scene)
        keep.geom = {k: full.geom[k] for k in ("radii", "blend_rec")}
        keep.binning = {"point_list": full.binning["point_list"], "ranges": full.binning["ranges"],
                        "scratch": self._tiny_scratch}         # the blend kernels use its first 256 bytes (unit counter)
        nbytes = sum(t.numel() * t.element_size() for t in list(keep.geom.values()) + [keep.binning["point_list"],
                                                                                       keep.binning["ranges"]])
        if self._geom_cache_used + nbytes > self.geometry_cache_bytes:
            return False
        self._geom_cache[view_id] = keep
        self._geom_cache_used += nbytes
        return True

    def sync_params(self) -> None:
        """Make every rank's logits / codebooks current (N > 1 with the deferred all-gather; no-op otherwise)."""
        if self.defer_gather and self.exchange is not None:
            self.exchange.all_gather_all()
            self.exchange.finish()

    def settings(self, cam: Dict) -> GaussianRasterizationSettings:
        d = self.dev
        return GaussianRasterizationSettings(
            image_height=int(cam["image_height"]), image_width=int(cam["image_width"]),
            tanfovx=float(cam["tanfovx"]), tanfovy=float(cam["tanfovy"]), bg=self.bg, scale_modifier=1.0,
            viewmatrix=cam["world_view_transform"].to(d, non_blocking=True),
            projmatrix=cam["full_proj_transform"].to(d, non_blocking=True),
            sh_degree=self.sh_degree, campos=cam["camera_center"].to(d, non_blocking=True), prefiltered=False, debug=False,
            include_feature=True, quick_render=False)

    def calibrate(self, cams, stride: int = 1) -> int:
        """Size the instance buffers from the largest of ``cams[::stride]`` (exact, synchronising frames) instead of
        from whatever the first few iterations happen to see."""
        from .rasterizer import rasterize_forward
        peak = 0
        with torch.no_grad():
            for cam in list(cams)[::max(1, stride)]:
                st = rasterize_forward(self.settings(cam), self.xyz, self.opacity, shs=self.shs, scales=self.scales,
                                       rotations=self.rotations, language_feature_precomp=self.logits.detach(),
                                       capacity=None, render=False)
                peak = max(peak, st.num_rendered)
        self._max_seen = max(self._max_seen, peak)
        self.capacity = int(self._max_seen * self.headroom) + 4096
        return self.capacity

    # ---- reference: compute_layer_feature_map (scene/gaussian_model.py:533-543) ----
    def layer_feature_map(self, wmap: torch.Tensor, layer_idx: int) -> torch.Tensor:
        cs = self.codebook_size
        feat = None
        for i in range(layer_idx + 1):
            f = decode_feature_map(wmap[i * cs:(i + 1) * cs], self.codebooks[i:i + 1], normalize=False,
                                   precision="fp32")[0]
            feat = f if feat is None else f + feat.detach()
        return feat

    def layer_index(self) -> int:
        return min(int(self.iteration / 10000 * self.layers), self.layers - 1)            # train.py:155

    def loss(self, cam: Dict, gt, mask: Optional[torch.Tensor], gt_sumsq: Optional[torch.Tensor] = None,
             view_id=None, _row_grad: bool = False) -> torch.Tensor:
        """The iteration's loss; ``loss.backward()`` fills ``logits.grad`` / ``codebooks.grad`` through autograd.
        (``_row_grad``: step()'s own form, whose backward stops at the blend and leaves the coefficient gradient in
        row form for _run_backward -- not for callers that run ``backward()`` themselves.)"""
        rs = self.settings(cam)
        gather = self.defer_gather and self.exchange is not None
        if gather:
            self.exchange.all_gather_all()          # last iteration's parameter shards -> every rank (side stream)
        prepared = self._geom_cache.get(view_id) if (view_id is not None and self._geom_cache) else None
        if prepared is not None:
            self.geometry_cache_hits += 1
        elif self.split_prepare and self.capacity is not None:
            # geometry stage first: it reads no trainable parameter, so it runs while the gather is in flight
            prepared = self._geometry_stage(rs)
        if gather:
            self.exchange.finish()
        rasterizer = GaussianRasterizer(raster_settings=rs, capacity=self.capacity)
        rasterizer.prepared = prepared
        rasterizer.dense_coeff_grad = False         # the top-k mask follows (utils/vq_utils.py:20): support-only is exact
        zero = torch.zeros((1,), dtype=torch.float32, device=self.dev)
        if _row_grad and self.row_grad and torch.is_grad_enabled():
            # ... and the blend backward hands its gradient back in row form: the top-k backward is run by
            # _run_backward / _finish_update (after the exchange at N > 1), not by autograd
            with torch.no_grad():
                coeff, rows = lssparse.softmax_to_topk_soft_code_rows(self.logits, self.topk)
            self._rows_g.zero_()
            rows["grad_rows"] = self._rows_g
            rows["dense"] = coeff.detach()           # what the top-k backward needs besides the rows
            coeff.requires_grad_(True)
            rasterizer.sparse_rows = rows            # lives on the autograd graph: a forward nobody backpropagates pins nothing
        elif self.fused_rows and self.layers == 1 and self.codebook_size == 64 and self.topk <= 16:
            # one level: the top-k kernel also writes the rows the blend reads (no compaction / packing launches)
            coeff, rasterizer.sparse_rows = lssparse.softmax_to_topk_soft_code_rows(self.logits, self.topk)
        else:
            coeff = self.render_weights()
        _, wmap, _ = rasterizer(means3D=self.xyz, means2D=self.means2D, shs=self.shs, colors_precomp=None,
                                language_feature_precomp=coeff, language_feature_weights_quick=zero,
                                language_feature_indices=zero, opacities=self.opacity, scales=self.scales,
                                rotations=self.rotations, cov3D_precomp=None)
        self._last_state = rasterizer.last_state
        layer_idx = self.layer_index()
        cs = self.codebook_size
        # one layer: the weight map IS the first layer's -- slicing it would make autograd pad the gradient back with a
        # zero fill + a 195 MB copy (0.46 ms at 1008x756, seen in the N = 2 timeline)
        w0 = wmap if wmap.shape[0] == cs else wmap[:cs]
        if isinstance(gt, (tuple, list)):
            # ground truth as (seg [H, W] int, table [S, 512]) -- scene/cameras.py:59-96 before the expansion
            seg, table = gt
            if self.gram_loss and layer_idx == 0 and cs == 64 and self.use_cos and not self.use_l1:
                return decode_cos_loss_segments(w0, self.codebooks[0:1], seg, table, normalize=self.normalize)
            valid = (seg >= 0)
            gt = table[seg.clamp(min=0).long()].permute(2, 0, 1).contiguous() * valid.unsqueeze(0)
            mask = valid.unsqueeze(0).to(torch.float32)
        if (self.gram_loss and self.fused_loss and layer_idx == 0 and cs == 64 and self.use_cos and not self.use_l1
                and mask.numel() == wmap[0].numel()):
            # |gt_p|^2: given by the caller, cached under the caller's view id, or recomputed (one streaming pass).
            # Never cached by address: a loader's fresh gt tensor can reuse the storage of another view's.
            g2 = gt_sumsq
            if g2 is None and view_id is not None:
                g2 = self._g2_by_view.get(view_id)
                if g2 is None:
                    if len(self._g2_by_view) > 256:
                        self._g2_by_view.clear()
                    g2 = self._g2_by_view[view_id] = column_sumsq(gt)
            if g2 is None:
                g2 = column_sumsq(gt)
            return decode_cos_loss(w0, self.codebooks[0:1], gt, mask, normalize=self.normalize, gt_sumsq=g2)
        feat = self.layer_feature_map(wmap, layer_idx)
        if self.fused_loss and self.use_cos and not self.use_l1 and mask.numel() == feat[0].numel():
            return normalized_cos_loss(feat, gt, mask, normalize=self.normalize)
        if self.normalize:
            feat = feat / (feat.norm(dim=0, keepdim=True) + 1e-10)
        loss = 0
        if self.use_cos:
            loss = loss + cos_loss(feat * mask, gt * mask)
        if self.use_l1:
            loss = loss + l1_loss(feat * mask, gt * mask)
        return loss

    # ------------------------------------------------------------------------------------------------------------
    def _backward_and_update(self, loss: torch.Tensor) -> None:
        """backward -> (N > 1: gradient exchange in row blocks) -> Adam; leaves the gradient buffer ready for the
        next iteration."""
        use_sink = self._sink_ok
        cbs = self._cb_sink
        if cbs is not None:
            from . import loss as _loss
            _loss.set_codebook_grad_sink(cbs)
            try:
                self._run_backward(loss, use_sink)
            finally:
                _loss.set_codebook_grad_sink(None)
            cbs.compute()                    # enqueued AFTER the blend / top-k backward and the logits' exchange
            self._finish_update(cbs.join, use_sink)
        else:
            self._run_backward(loss, use_sink)
            self._finish_update(lambda: None, use_sink)

    def _run_backward(self, loss: torch.Tensor, use_sink: bool) -> None:
        n_log = self.n_log
        if self.row_grad:
            from .rasterizer import _RasterizeGaussians
            _RasterizeGaussians.last_rows = None
            loss.backward()                          # ends with the blend backward: self._rows_g holds this rank's rows
            rows, _RasterizeGaussians.last_rows = _RasterizeGaussians.last_rows, None
            if rows is None or rows.get("grad_rows") is not self._rows_g:
                raise RuntimeError("the blend backward did not hand back the row-form gradient of this trainer's loss()")
            if self.rows_exchange is not None:
                for i in range(len(self._chunk_rows)):
                    self.rows_exchange.reduce_scatter(i)
                self._row_ctx = (rows["dense"], rows)                    # until _finish_update
            else:
                lssparse.topk_backward_rows(self.logits, rows["dense"], rows, self._rows_g, self.logits.grad)
            return
        if use_sink:
            if self.collective == "zero1":
                on_chunk = lambda a, b, _s=self: _s.exchange.reduce_scatter(_s._block_of[a])
            elif self.collective == "allreduce":
                self.bucket.begin_overlap()
                on_chunk = self.bucket.reduce_slice
            else:
                on_chunk = lambda a, b: None
            self._block_of = {a: i for i, (a, _) in enumerate(self._blocks)}
            sink = lssparse.GradSink(self.logits.grad, self._chunk_rows, on_chunk)
            lssparse.set_grad_sink(sink)
            try:
                loss.backward()
            finally:
                lssparse.set_grad_sink(None)
            if sink.used_upto != n_log:
                raise RuntimeError("the top-k backward did not take the gradient sink")
        else:
            loss.backward()

    def _finish_update(self, join_codebook_grad, use_sink: bool) -> None:
        """Gradient exchange (N > 1) and Adam.  ``join_codebook_grad()`` makes the current stream wait for the deferred
        codebook gradient; it is called as late as possible -- after the logits' blocks have been exchanged / stepped."""
        opt = self.optimizer
        n_log = self.n_log
        nb = len(self._blocks)
        if self.collective == "zero1":
            ex = self.exchange
            if not use_sink and self._row_ctx is None:
                for i in range(nb - 1):
                    ex.reduce_scatter(i)
            defer = self.defer_gather
            row_ctx, self._row_ctx = self._row_ctx, None
            C = self.logits.shape[1]
            for i in range(nb - 1):                                      # the logits' row blocks
                lo, hi = ex.shard_range(i)
                if row_ctx is not None:
                    # this rank's rows of block i, summed over the ranks -> d logits of those rows only
                    rex = self.rows_exchange
                    rex.wait_grad(i)
                    lssparse.topk_backward_rows(self.logits, row_ctx[0], row_ctx[1],
                                                rex.shard_grad[i].view(-1, self._row_stride),
                                                ex.shard_grad[i].view(-1, C), row_range=(lo // C, hi // C))
                else:
                    ex.wait_grad(i)
                opt.step_range(lo, hi, grad=ex.shard_grad[i])
                if not defer:
                    ex.all_gather_params(i)
            join_codebook_grad()
            ex.reduce_scatter(nb - 1)                                    # codebooks (+ pad)
            ex.wait_grad(nb - 1)
            lo, hi = ex.shard_range(nb - 1)
            opt.step_range(lo, hi, grad=ex.shard_grad[nb - 1])
            if not defer:
                ex.all_gather_params(nb - 1)
            if self.keep_last_grad:
                self.last_grad = ex.full_reduced_grad()
            if not defer:
                ex.finish()
        elif self.collective == "allreduce":
            bk = self.bucket
            if not use_sink:
                bk.begin_overlap()
                for a, b in self._blocks[:-1]:
                    bk.reduce_slice(a, b)
            join_codebook_grad()
            bk.reduce_slice(*self._blocks[-1])
            for i, (a, b) in enumerate(self._blocks):
                bk.wait_slice(i)
                opt.step_range(a, b)
            bk.finish_overlap(self.flat_g.numel())
            if self.keep_last_grad:
                self.last_grad = self.flat_g.clone()
        else:
            if use_sink and not self.keep_last_grad:
                opt.step_range(0, n_log)                                 # fully rewritten by the next top-k backward
                join_codebook_grad()
                opt.step_range(n_log, self.flat_p.numel(), zero_grad=True)
            else:
                join_codebook_grad()
                if self.keep_last_grad:
                    self.last_grad = self.flat_g.clone()
                if use_sink:
                    opt.step_range(0, n_log)
                    opt.step_range(n_log, self.flat_p.numel(), zero_grad=True)
                else:
                    opt.step_range(0, self.flat_p.numel(), zero_grad=True)
            return
        # N > 1: what autograd accumulates into must be zero again (the sink overwrites the logits part)
        if use_sink:
            self.flat_g[n_log:].zero_()
        else:
            self.flat_g.zero_()

    def _iteration(self, cam, gt, mask, gt_sumsq, view_id) -> torch.Tensor:
        """forward + backward + update; the caller has advanced the optimizer's step count (begin_step)."""
        loss = self.loss(cam, gt, mask, gt_sumsq=gt_sumsq, view_id=view_id, _row_grad=True)
        self._backward_and_update(loss)
        st = self._last_state
        if st is not None and self.capacity is not None:
            # sticky on the device: [1] |= overflow of this frame
            self._sticky[0:1].bitwise_or_(st.status[1:2])
        return loss.detach()

    def step(self, cam: Dict, gt, mask: Optional[torch.Tensor] = None, gt_sumsq: Optional[torch.Tensor] = None,
             view_id=None) -> torch.Tensor:
        """One iteration; returns the (detached, on-device) loss.  ``view_id`` (any hashable) lets the trainer cache
        per-view constants (|gt|^2) and, with ``use_graph``, the captured iteration of that view; the tensors passed
        for a given ``view_id`` must then be the same (resident) tensors on every call."""
        self.iteration += 1
        calibrated = self.capacity is not None
        # the first iteration always runs eagerly: libraries create handles / workspaces on first use (cuBLAS, NCCL
        # communicators), which is not allowed under stream capture
        if self.use_graph and calibrated and view_id is not None and gt_sumsq is None and self._eager_iters > 0:
            key = (view_id, self.layer_index())
            entry = self._graphs.get(key)
            if entry is None:
                if view_id not in self._g2_by_view and not isinstance(gt, (tuple, list)) and self.gram_loss:
                    self._g2_by_view[view_id] = column_sumsq(gt)         # outside the capture
                self.optimizer.device_scalars = True                     # from now on the kernels read them from the device
                torch.cuda.synchronize(self.dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, pool=self._graph_pool):
                    out = self._iteration(cam, gt, mask, None, view_id)   # captured, not run
                if self._graph_pool is None:
                    self._graph_pool = g.pool()
                entry = self._graphs[key] = (g, out)
            g, out = entry
            self.optimizer.begin_step()                                  # step count + device-side scalars, outside the graph
            g.replay()
            loss = out
        else:
            self.optimizer.begin_step()
            loss = self._iteration(cam, gt, mask, gt_sumsq, view_id)
            self._eager_iters += 1
        st = self._last_state
        if not calibrated and st is not None:
            self._max_seen = max(self._max_seen, st.num_rendered)
            if self.iteration >= self.calib_steps:
                self.capacity = int(self._max_seen * self.headroom) + 4096
        elif self.check_every > 0 and self.iteration % self.check_every == 0:
            if int(self._sticky[0].item()):                              # one small D2H read every check_every steps
                # some frame since the last check was rendered from a truncated instance list: grow and go on
                self.overflow_events += 1
                self._sticky.zero_()
                self.capacity = int(self.capacity * 1.5) + 4096
                self._graphs.clear()
                self._geom_cache.clear()                                 # cached lists were built under the old capacity
                self._geom_cache_used = 0
                import warnings
                warnings.warn(f"instance capacity overflowed; grown to {self.capacity} (the affected iterations used "
                              "truncated instance lists)")
        self._last_state = None
        return loss


def synthetic_gt_segments(H: int, W: int, feat_dim: int = 512, seed: int = 2, device="cuda"):
    """The same ground truth as ``synthetic_gt`` in the loader's form: (seg [H, W] int32, table [256, feat_dim])."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    table = torch.randn((256, feat_dim), generator=g)
    table = table / table.norm(dim=1, keepdim=True)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    seg = ((xs // 32 + 7 * (ys // 32)) % 256).to(torch.int32)
    return seg.to(device), table.to(device)


def synthetic_gt(H: int, W: int, feat_dim: int = 512, seed: int = 2, device="cuda"):
    """piecewise-constant ground truth, seg = (x//32 + 7*(y//32)) % 256, rows of a unit-norm table;
    mask all true.  -> (gt [feat_dim, H, W], mask [1, H, W])"""
    g = torch.Generator(device="cpu").manual_seed(seed)
    table = torch.randn((256, feat_dim), generator=g)
    table = table / table.norm(dim=1, keepdim=True)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    seg = (xs // 32 + 7 * (ys // 32)) % 256
    gt = table.to(device)[seg.to(device)].permute(2, 0, 1).contiguous()
    mask = torch.ones((1, H, W), dtype=torch.float32, device=device)
    return gt, mask
