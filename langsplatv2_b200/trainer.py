"""Feature-stage training step on the B200 op (the "train iter/s" half of the metric).

Mirrors ``train.py --include_feature`` of the reference, one iteration per ``step()``:

    render (include_feature)          /root/reference/train.py:147, gaussian_renderer/__init__.py:95-98,108-119
      logits -> softmax/top-k/renorm  scene/gaussian_model.py:510-518, utils/vq_utils.py:9-24
    decode one codebook layer         scene/gaussian_model.py:533-543 (layer_num = 1 in the paper, train.py:154)
    normalise, cosine (+ L1) loss     train.py:157-167, utils/loss_utils.py:18-25
    backward                          train.py:173
    Adam on (logits, codebooks)       scene/gaussian_model.py:235-236,255 ; train.py:261-263

Geometry is frozen (scene/gaussian_model.py:238-243).  Data parallel: each rank renders its own view and the
gradients of (logits, codebooks) are summed over the ranks in ONE flat bucket (``dist.GradBucket``; reduced in 4 row
slices pipelined with the top-k backward and per-slice Adam steps), SUM as in the reference's ``--accum_iter``
(train.py:261-263,348).  The hot kernels (top-k fwd/bwd, preprocess, binning, blend fwd/bwd,
decode fwd/bwd) are liblsv2's; the loss expression and Adam are PyTorch's, as they are in the reference.
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from . import dist as lsdist
from .decode import decode_feature_map
from .loss import column_sumsq, decode_cos_loss, decode_cos_loss_segments, normalized_cos_loss
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
    """Resident scene + the two trainable tensors + Adam; ``step(cam, gt, mask)`` is one train.py iteration."""

    def __init__(self, scene: Dict, topk: int = 4, lr: float = 0.0025, device="cuda", normalize: bool = True,
                 use_cos: bool = True, use_l1: bool = False, layers: Optional[int] = None, fused_adam: bool = True,
                 fused_loss: bool = True, gram_loss: bool = True):
        dev = torch.device(device)
        f32 = lambda t: t.to(dev, torch.float32).contiguous()
        self.dev = dev
        self.topk = topk
        self.normalize, self.use_cos, self.use_l1 = normalize, use_cos, use_l1
        self.fused_loss = fused_loss        # cos loss (+ the normalisation) as two kernels instead of torch's ~28 passes
        self.gram_loss = gram_loss          # ... and, for a single layer, without forming the [512, H, W] feature map
        self._g2_cache = {}                 # |gt_p|^2 per ground-truth tensor (constant across iterations of a view)
        self.xyz, self.opacity = f32(scene["xyz"]), f32(scene["opacity"])
        self.scales, self.rotations = f32(scene["scaling"]), f32(scene["rotation"])
        self.shs = f32(scene["features"])
        self.sh_degree = int(scene.get("sh_degree", 3))
        cb = f32(scene["codebooks"])
        self.layers = int(cb.shape[0] if layers is None else layers)
        self.codebook_size = int(cb.shape[1])
        self.logits = torch.nn.Parameter(f32(scene["logits"])[:, : self.layers * self.codebook_size].clone())
        self.codebooks = torch.nn.Parameter(cb[: self.layers].clone())
        self.bg = torch.zeros(3, dtype=torch.float32, device=dev)
        self.means2D = torch.zeros_like(self.xyz)                      # grad sink of the reference API; unused here
        # the reference: torch.optim.Adam(l, lr=0.0, eps=1e-15) with the group lr = language_feature_lr
        self.optimizer = torch.optim.Adam([{"params": [self.logits, self.codebooks], "lr": lr}], lr=0.0, eps=1e-15,
                                          fused=fused_adam and dev.type == "cuda")
        self.bucket = None
        self.overlap_chunks = int(os.environ.get("LSV2_OVERLAP_CHUNKS", "4"))
        self._chunk_rows = None
        self._chunk_opts = None
        self.keep_last_grad = False
        self.last_grad = None
        if lsdist.world_info()[1] > 1:
            self.bucket = lsdist.GradBucket([self.logits, self.codebooks])
            self.bucket.attach()                    # grads are written into the flat all-reduce buffer directly
            if self.layers == 1 and dev.type == "cuda":
                # row blocks of the logits: block i's gradient leaves the top-k backward, is all-reduced on a side
                # stream, and gets its own (elementwise-identical) Adam step while block i + 1 is still in flight
                n = self.logits.shape[0]
                rows = -(-(-(-n // self.overlap_chunks)) // 1024) * 1024
                self._chunk_rows = [(a, min(n, a + rows)) for a in range(0, n, rows)]
                gview = self.bucket.views[0]
                self._chunk_opts = []
                for a, b in self._chunk_rows:
                    p = self.logits.detach()[a:b]                   # leaf view of the same storage
                    p.grad = gview[a:b]
                    self._chunk_opts.append(torch.optim.Adam([{"params": [p], "lr": lr}], lr=0.0, eps=1e-15,
                                                             fused=fused_adam))
                cb = self.codebooks.detach()
                cb.grad = self.bucket.views[1]
                self._cb_opt = torch.optim.Adam([{"params": [cb], "lr": lr}], lr=0.0, eps=1e-15, fused=fused_adam)
        self.iteration = 0
        # instance capacity: the first `calib_steps` iterations read num_rendered back (the reference's one host sync
        # per frame); after that the frame runs with a bounded instance buffer and no host sync
        self.capacity: Optional[int] = None
        self.calib_steps = 3
        self._max_seen = 0
        self._last_state = None

    # ---- reference: GaussianModel.get_render_weights (scene/gaussian_model.py:510-518) ----
    def render_weights(self) -> torch.Tensor:
        cs = self.codebook_size
        w = [softmax_to_topk_soft_code(self.logits[:, i * cs:(i + 1) * cs], self.topk) for i in range(self.layers)]
        return w[0] if len(w) == 1 else torch.cat(w, dim=-1)

    def settings(self, cam: Dict) -> GaussianRasterizationSettings:
        d = self.dev
        return GaussianRasterizationSettings(
            image_height=int(cam["image_height"]), image_width=int(cam["image_width"]),
            tanfovx=float(cam["tanfovx"]), tanfovy=float(cam["tanfovy"]), bg=self.bg, scale_modifier=1.0,
            viewmatrix=cam["world_view_transform"].to(d, non_blocking=True),
            projmatrix=cam["full_proj_transform"].to(d, non_blocking=True),
            sh_degree=self.sh_degree, campos=cam["camera_center"].to(d, non_blocking=True), prefiltered=False, debug=False,
            include_feature=True, quick_render=False)

    # ---- reference: compute_layer_feature_map (scene/gaussian_model.py:533-543) ----
    def layer_feature_map(self, wmap: torch.Tensor, layer_idx: int) -> torch.Tensor:
        cs = self.codebook_size
        feat = None
        for i in range(layer_idx + 1):
            f = decode_feature_map(wmap[i * cs:(i + 1) * cs], self.codebooks[i:i + 1], normalize=False,
                                   precision="fp32")[0]
            feat = f if feat is None else f + feat.detach()
        return feat

    def loss(self, cam: Dict, gt: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
        rasterizer = GaussianRasterizer(raster_settings=self.settings(cam), capacity=self.capacity)
        zero = torch.zeros((1,), dtype=torch.float32, device=self.dev)
        _, wmap, _ = rasterizer(means3D=self.xyz, means2D=self.means2D, shs=self.shs, colors_precomp=None,
                                language_feature_precomp=self.render_weights(), language_feature_weights_quick=zero,
                                language_feature_indices=zero, opacities=self.opacity, scales=self.scales,
                                rotations=self.rotations, cov3D_precomp=None)
        self._last_state = rasterizer.last_state
        layer_idx = min(int(self.iteration / 10000 * self.layers), self.layers - 1)       # train.py:155
        cs = self.codebook_size
        if isinstance(gt, (tuple, list)):
            # ground truth as (seg [H, W] int, table [S, 512]) -- scene/cameras.py:59-96 before the expansion
            seg, table = gt
            if self.gram_loss and layer_idx == 0 and cs == 64 and self.use_cos and not self.use_l1:
                return decode_cos_loss_segments(wmap[:cs], self.codebooks[0:1], seg, table, normalize=self.normalize)
            valid = (seg >= 0)
            gt = table[seg.clamp(min=0).long()].permute(2, 0, 1).contiguous() * valid.unsqueeze(0)
            mask = valid.unsqueeze(0).to(torch.float32)
        if (self.gram_loss and self.fused_loss and layer_idx == 0 and cs == 64 and self.use_cos and not self.use_l1
                and mask.numel() == wmap[0].numel()):
            key = (gt.data_ptr(), tuple(gt.shape))
            g2 = self._g2_cache.get(key)
            if g2 is None:
                if len(self._g2_cache) > 64:
                    self._g2_cache.clear()
                g2 = self._g2_cache[key] = column_sumsq(gt)
            return decode_cos_loss(wmap[:cs], self.codebooks[0:1], gt, mask, normalize=self.normalize, gt_sumsq=g2)
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

    def step(self, cam: Dict, gt: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
        """One iteration; returns the (detached, on-device) loss."""
        self.iteration += 1
        loss = self.loss(cam, gt, mask)
        if self._chunk_opts is not None and self.overlap_chunks > 1:
            # the flat bucket [d logits || d codebooks] is reduced in slices: d logits leaves the top-k backward in
            # row blocks, block i's all-reduce runs under block i + 1's kernel and under block i - 1's Adam
            self.bucket.begin_overlap()
            sink = lssparse.GradSink(self.bucket.views[0], self._chunk_rows, self.bucket.reduce_slice)
            lssparse.set_grad_sink(sink)
            try:
                loss.backward()
            finally:
                lssparse.set_grad_sink(None)
            if sink.used_upto == self.logits.numel():
                self.bucket.reduce_slice(sink.used_upto, self.bucket.flat.numel())      # the codebook part
                for i, opt in enumerate(self._chunk_opts):
                    self.bucket.wait_slice(i)
                    opt.step()
                self.bucket.wait_slice(len(self._chunk_opts))
                self._cb_opt.step()
                self.bucket.finish_overlap(self.bucket.flat.numel())
            else:                                   # the sink was not taken (shape mismatch): classic path
                self.bucket.finish_overlap(0)
                for opt in self._chunk_opts:
                    opt.step()
                self._cb_opt.step()
        else:
            loss.backward()
            if self.bucket is not None:
                self.bucket.sync()                  # ONE all-reduce(SUM) of [d logits || d codebooks]
            if self._chunk_opts is not None:
                for opt in self._chunk_opts:
                    opt.step()
                self._cb_opt.step()
            else:
                self.optimizer.step()
        if self.bucket is not None:
            if self.keep_last_grad:
                self.last_grad = self.bucket.flat.clone()           # the all-reduced gradient of this step (tests)
            self.bucket.zero()
        else:
            self.optimizer.zero_grad(set_to_none=True)
        st = self._last_state
        if st is not None:
            if self.capacity is None:
                self._max_seen = max(self._max_seen, st.num_rendered)
                if self.iteration >= self.calib_steps:
                    self.capacity = int(self._max_seen * 1.5) + 4096
            elif self.iteration % 64 == 0 and st.overflowed:       # rare check (one small D2H read)
                raise RuntimeError(f"instance capacity {self.capacity} exceeded; raise FeatureTrainer.capacity")
        self._last_state = None
        return loss.detach()


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
