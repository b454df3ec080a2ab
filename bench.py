#!/usr/bin/env python
"""bench.py -- 512-d feature-map FPS of the LangSplatV2 rasterization hot path on N B200s.

A "step" is one frame: camera -> preprocess -> binning -> blend (quick mode, 3 levels x top-4 of 64)
-> codebook decode -> L2-normalised [3, 512, H, W] feature map.  Default workload: 1M synthetic Gaussians
at 1920x1080 (the configuration BASELINE.json's metric is quoted on); ``--workload c2`` is 1008x756.
View-sharded across ranks (weak scaling: every rank renders --steps frames of its own views; no data-path
collective).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload 1080p|c2|4k|small] [--gaussians N]
    python bench.py --workload c1|c4 ...      # BASELINE.json configs 1 / 4: forward + backward iterations

``--impl reference`` times the reference's CPU implementation of the same path -- here the CPU restatement
in oracle/ (the reference's rasterizer source is absent from the mounted scrape) -- on the
host cores, on a bounded tile sample of the same frame (the extrapolation factor is stated in the line).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "512-d feature-map FPS (1M Gaussians, 1080p)"
UNIT = "frames/s"
WORKLOADS = {
    "1080p": dict(N=1_000_000, W=1920, H=1080, name="1M synthetic Gaussians, 1920x1080, 3 levels x top-4 of 64, 512-d normalised fp32 feature map (quick render + decode)"),
    "c2": dict(N=1_000_000, W=1008, H=756, name="1M synthetic Gaussians, 1008x756, 3 levels x top-4 of 64, 512-d normalised fp32 feature map (quick render + decode)"),
    "4k": dict(N=1_000_000, W=3840, H=2160, name="1M synthetic Gaussians, 3840x2160, 3 levels x top-4 of 64, 512-d normalised fp32 feature map (BASELINE.json config 5)"),
    "small": dict(N=100_000, W=640, H=360, name="100k synthetic Gaussians, 640x360, 3 levels (debug size)"),
    # forward + backward workloads (one FeatureTrainer iteration per step)
    "c1": dict(N=50_000, W=256, H=256, fwd_bwd=True, name="BASELINE.json config 1: 50k synthetic Gaussians, 256x256, top-4 of 64, 512-d decode, forward + backward + Adam"),
    "c4": dict(N=4_000_000, W=1600, H=1066, fwd_bwd=True, name="BASELINE.json config 4: 4M synthetic Gaussians, 1600x1066, top-4 of 64, 512-d decode, forward + backward + Adam"),
}
VIEWS = 64
LEVELS, CODES, D, TOPK = 3, 64, 512, 4


def ncu_traffic(kernel_substr: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, read at bench time from the
    newest `ncu --set full` summary committed under profiles/ (profiles/summarize.py output) -- never a literal.
    -> (bytes or None, "file: kernel line" or None)."""
    import glob
    import re
    best = None
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r0[2-9]*_ncu_*.txt"))):
        name, rd, wr = None, None, None
        for ln in open(path, errors="replace"):
            if ln.startswith("## "):
                if name and rd is not None and wr is not None and kernel_substr in name:
                    best = (rd + wr, f"{os.path.basename(path)}: {name}")
                name, rd, wr = ln[3:].strip(), None, None
            m = re.match(r"\s*dram__bytes_(read|write)\.sum\s+([0-9.eE+-]+)\s+(\w+)", ln)
            if m:
                v = float(m.group(2)) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(m.group(3), 1)
                if m.group(1) == "read":
                    rd = v
                else:
                    wr = v
        if name and rd is not None and wr is not None and kernel_substr in name:
            best = (rd + wr, f"{os.path.basename(path)}: {name}")
    return best if best else (None, None)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self, wait_first: float = 10.0):
        """Starts nvidia-smi and waits until its first line is out (its start-up takes longer than a short timed
        region), so that every later line is a sample under whatever the GPU is doing then."""
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(self.gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None
            return
        t0 = time.perf_counter()
        while self.lines() == 0 and time.perf_counter() - t0 < wait_first and self.p.poll() is None:
            time.sleep(0.02)
        self.skip = self.lines()          # samples before the load starts are not counted

    def lines(self) -> int:
        try:
            with open(self.f.name) as fh:
                return sum(1 for _ in fh)
        except OSError:
            return 0

    def samples_under_load(self) -> int:
        return max(0, self.lines() - getattr(self, "skip", 0))

    def keep_loaded(self, step_fn, want: int = 3, limit_s: float = 3.0) -> None:
        """A timed region shorter than a few sampling periods: keep the SAME load running, untimed, until ``want``
        samples exist (the clocks line is about the load, not about the stopwatch)."""
        if self.p is None:
            return
        t0 = time.perf_counter()
        while self.samples_under_load() < want and time.perf_counter() - t0 < limit_s:
            step_fn()

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines()[getattr(self, "skip", 0):]:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return out


def make_inputs(wl):
    from langsplatv2_b200 import synth
    scene = synth.make_scene(wl["N"], seed=0, levels=LEVELS, codebook_size=CODES, feat_dim=D, sh_degree=3)
    cams = [synth.ring_camera(v, VIEWS, wl["W"], wl["H"]) for v in range(VIEWS)]
    return scene, cams


# ----------------------------------------------------------------------------------------------
# CPU arm: the checker's restatement of the path, timed on the host cores on a bounded tile sample
# ----------------------------------------------------------------------------------------------
def cpu_frame_time(scene, cam, wl, budget_s: float):
    """-> (estimated seconds per full frame, description of the sample, threads).
    preprocess + binning run on the whole frame; blend + decode run on every `stride`-th tile and are
    scaled by `stride` (tiles are independent, so the extrapolation is linear in the tile count)."""
    import oracle
    W, H = wl["W"], wl["H"]
    logits = scene["logits"].numpy()
    ws, idx = [], []
    for l in range(LEVELS):
        w, i = oracle.weights_and_indices(logits[:, l * CODES:(l + 1) * CODES], TOPK)
        ws.append(w)
        idx.append(i + l * CODES)
    wq, iq = np.concatenate(ws, 1), np.concatenate(idx, 1)
    osc = dict(means3d=scene["xyz"].numpy(), opacities=scene["opacity"].numpy(), scales=scene["scaling"].numpy(),
               rotations=scene["rotation"].numpy(), shs=scene["features"].numpy(), sh_degree=scene["sh_degree"])
    t0 = time.perf_counter()
    pre = oracle.preprocess(osc["means3d"], osc["opacities"], cam["world_view_transform"].numpy(),
                            cam["full_proj_transform"].numpy(), cam["camera_center"].numpy(), W, H,
                            cam["tanfovx"], cam["tanfovy"], scales=osc["scales"], rotations=osc["rotations"],
                            shs=osc["shs"], sh_degree=osc["sh_degree"])
    binned = oracle.bin_instances(pre, W, H)
    t_front = time.perf_counter() - t0
    n_tiles = ((W + 15) // 16) * ((H + 15) // 16)
    # ~25 ms of CPU work per tile (blend + 3-level decode of 256 px) on 8 cores is the planning figure
    stride = max(1, int(math.ceil(n_tiles * 0.025 / max(budget_s - t_front, 1.0))))
    t0 = time.perf_counter()
    fwd = oracle.render_forward(pre, binned, W, H, (0, 0, 0), mode=2, L=LEVELS * CODES, weights_quick=wq,
                                indices_quick=iq, tile_stride=stride, tile_phase=0)
    mask = oracle.tile_pixel_mask(W, H, stride, 0)
    wsel = np.ascontiguousarray(fwd["wmap"][:, mask])
    feat = oracle.decode(wsel, scene["codebooks"].numpy(), normalize=True)
    t_back = time.perf_counter() - t0
    assert feat.shape[0] == LEVELS
    est = t_front + t_back * stride
    sample = (f"preprocess+binning of the full frame ({t_front:.2f} s) + blend+decode of every {stride}-th 16x16 tile "
              f"({int(mask.sum())} of {W * H} px, {t_back:.2f} s) scaled x{stride}")
    return est, sample, oracle.num_threads()


def reference_cuda_rasterizer_status() -> str:
    """the reference's own CUDA rasterizer would be timed here if a driver had installed it
    under baseline/_ref (package diff_gaussian_rasterization with its compiled _C).  Its source is absent from the
    scrape (.SUBMODULES.json: submodules/efficient-langsplat-rasterization missing), so there is nothing to install."""
    d = os.path.join(ROOT, "baseline", "_ref", "diff_gaussian_rasterization")
    if os.path.isdir(d) and any(f.startswith("_C") for f in os.listdir(d)):
        return "present under baseline/_ref (not timed by this harness version)"
    return "unavailable on this box: the submodule source is absent from the scrape, baseline/_ref holds no build of it"


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # torchrun exports OMP_NUM_THREADS=1: the CPU arm is one process on rank 0 and takes every host core
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    if wl.get("fwd_bwd"):
        return run_reference_fwd_bwd(args, wl)
    scene, cams = make_inputs(wl)
    budget = max(5.0, min(40.0, 150.0 / max(args.steps + args.warmup, 1)))
    times, sample, threads = [], "", 1
    for s in range(args.warmup + args.steps):
        est, sample, threads = cpu_frame_time(scene, cams[s % VIEWS], wl, budget)
        if s >= args.warmup:
            times.append(est)
    sec = float(np.mean(times))
    val = args.gpus * 0 + 1.0 / sec
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": wl["name"], "views": VIEWS,
                       "note": "CPU restatement (oracle/) on all host cores of rank 0; blend + decode timed on a tile sample and "
                               "extrapolated linearly in the tile count (factor in cpu_baseline.sample)",
                       "reference_cuda_rasterizer": reference_cuda_rasterizer_status()},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def cpu_fwd_bwd_time(wl, reps: int = 1):
    """BASELINE.json config 1 verbatim: the pure-PyTorch evaluation of the rasterizer math (oracle/torch_ref.py), forward
    + backward of one training iteration (top-k soft code -> render -> decode -> normalise -> cos loss), in FULL (no
    tile sampling), on the host cores.  -> seconds per iteration."""
    from langsplatv2_b200 import synth
    from oracle import torch_ref
    scene = synth.make_scene(wl["N"], seed=0, levels=1, codebook_size=CODES, feat_dim=D, sh_degree=3)
    cam = synth.ring_camera(0, VIEWS, wl["W"], wl["H"])
    W, H = wl["W"], wl["H"]
    g = torch.Generator().manual_seed(2)
    table = torch.nn.functional.normalize(torch.randn(256, D, generator=g), dim=1)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    gt = table[(xs // 32 + 7 * (ys // 32)) % 256].permute(2, 0, 1).contiguous()
    times = []
    for _ in range(reps):
        t0 = time.perf_counter()
        logits = scene["logits"][:, :CODES].clone().requires_grad_(True)
        cb = scene["codebooks"][:1].clone().requires_grad_(True)
        pre = torch_ref.preprocess(scene["xyz"], scene["opacity"], scene["scaling"], scene["rotation"],
                                   cam["world_view_transform"], cam["full_proj_transform"], cam["camera_center"], W, H,
                                   cam["tanfovx"], cam["tanfovy"], shs=scene["features"], sh_degree=3)
        binned = torch_ref.bin_instances(pre, W, H)
        dense = torch_ref.topk_soft_code(logits, TOPK)
        out = torch_ref.render(pre, binned, W, H, torch.zeros(3), coeff=dense)
        feat = torch_ref.decode(out["wmap"], cb, normalize=False)[0]
        feat = feat / (feat.norm(dim=0, keepdim=True) + 1e-10)
        loss = 1 - torch.nn.functional.cosine_similarity(feat, gt, dim=0).mean()
        loss.backward()
        times.append(time.perf_counter() - t0)
    return float(np.mean(times)), torch.get_num_threads()


def run_reference_fwd_bwd(args, wl):
    steps = max(1, min(args.steps, 3))
    if wl["N"] > 200_000:
        line = {"impl": "reference", "unavailable": "the pure-PyTorch CPU evaluation is only run in full at BASELINE.json config 1 (50k, 256x256); "
                                                     "at this size one iteration takes hours"}
        print(json.dumps(line), flush=True)
        return
    sec, threads = cpu_fwd_bwd_time(wl, reps=steps)
    val = 1.0 / sec
    line = {"metric": "forward + backward iterations/s", "value": val, "unit": "iter/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": 0, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": wl["name"], "note": "oracle/torch_ref.py (pure PyTorch, CPU, autograd), whole frame, no sampling"},
            "cpu_baseline": {"value": val, "unit": "iter/s", "cores": threads, "kind": "port", "sample": "full iteration, no sampling"},
            "e2e": {"value": val, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def kernels_per_frame(W, H, fused: bool) -> int:
    """liblsv2 kernels launched per frame (torch's own fill kernels are not counted)."""
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    tb = max(1, math.ceil(math.log2(tiles)))
    # preprocess, scan | depth keys, histogram, 4 depth passes over the Gaussians, gather, scan | duplicate, tile
    # histogram, ranges, tile passes over the instances
    n = 1 + 1 + (1 + 1 + (31 + 7) // 8 + 1 + 1) + (1 + 1 + 1 + (tb + 7) // 8)
    n += 3                                  # codebook image + Gram image + fused blend/decode | blend + codebook prep + decode
    return n


TRAIN_WL = dict(N=1_000_000, W=1008, H=756,
                name="train.py --include_feature step shape: frozen 1M-Gaussian geometry, 1008x756, 1 codebook layer "
                     "(64 x 512), top-4, normalise + cos loss, Adam on logits [1M,64] + codebook")


def _timed_iterations(tr, cams, gts, my_views, steps, dev, world, barrier, seg=False):
    """`steps` iterations, one CUDA event per iteration on the launching stream -> per-iteration ms (max over ranks per
    iteration at N > 1 is not available without a sync, so the run's total is the max over ranks and the
    distribution is rank 0's)."""
    barrier()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    loss = None
    for s in range(steps):
        v = my_views[s % len(my_views)]
        if seg:
            loss = tr.step(cams[v], gts[v], None, view_id=("seg", v))
        else:
            loss = tr.step(cams[v], *gts[v], view_id=v)
        evs[s + 1].record()
    barrier()
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    t = torch.tensor([evs[0].elapsed_time(evs[steps])], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), per, float(loss)


def run_train(args, dev, rank, world, scene, barrier, wl=None):
    """Second half of the metric ("train iter/s"): one iteration = softmax/top-k -> render (include_feature) ->
    decode -> normalise -> cos loss -> backward -> (N>1: gradient exchange) -> Adam.
    Weak scaling: every rank renders its own view each iteration (the reference's --accum_iter N)."""
    from langsplatv2_b200 import synth
    from langsplatv2_b200.trainer import FeatureTrainer, synthetic_gt, synthetic_gt_segments
    default_wl = wl is None
    wl = wl or TRAIN_WL
    W, H = wl["W"], wl["H"]
    steps = max(1, args.train_steps)
    cams = [synth.ring_camera(v, VIEWS, W, H) for v in range(VIEWS)]
    my_views = [v for v in range(VIEWS) if v % world == rank][:4]
    cams = {v: {k: (x.to(dev) if torch.is_tensor(x) else x) for k, x in cams[v].items()} for v in my_views}
    gts = {v: synthetic_gt(H, W, device=dev, seed=2 + v) for v in my_views}
    use_graph = not args.no_graph
    tr = FeatureTrainer(scene, topk=TOPK, layers=1, device=dev, use_graph=use_graph)
    tr.calibrate([cams[v] for v in my_views])          # exact instance counts of the views this rank trains on
    n_warm = (2 * len(my_views) + 2) if world == 1 else 12
    for v in my_views:                                  # one eager iteration per view (lazy initialisations, allocator)
        tr.step(cams[v], *gts[v])
    graph_note = ("CUDA graph per view (captured after calibration%s)" % (", NCCL exchange inside the graph" if world > 1 else "")) if use_graph else "eager launches"
    try:
        for s in range(n_warm):
            v = my_views[s % len(my_views)]
            tr.step(cams[v], *gts[v], view_id=v)
    except Exception as e:                              # capture refused: fall back to eager launches, say so
        if not use_graph:
            raise
        torch.cuda.synchronize(dev)
        tr = FeatureTrainer(scene, topk=TOPK, layers=1, device=dev, use_graph=False)
        tr.calibrate([cams[v] for v in my_views])
        graph_note = "eager launches (graph capture failed: %s)" % str(e)[:120]
        for s in range(n_warm):
            v = my_views[s % len(my_views)]
            tr.step(cams[v], *gts[v], view_id=v)
    ms, per, loss = _timed_iterations(tr, cams, gts, my_views, steps, dev, world, barrier)
    # the same iterations with the ground truth in the reference loader's own form (segmentation map + per-segment
    # table, scene/cameras.py:59-96) instead of the expanded [512, H, W] tensor: loss.decode_cos_loss_segments
    segs = {v: synthetic_gt_segments(H, W, device=dev, seed=2 + v) for v in my_views}
    for s in range(2 * len(my_views)):
        v = my_views[s % len(my_views)]
        tr.step(cams[v], segs[v], None, view_id=("seg", v))
    ms_seg, per_seg, loss_seg = _timed_iterations(tr, cams, segs, my_views, steps, dev, world, barrier, seg=True)
    pct = lambda xs, q: float(np.percentile(np.asarray(xs), q))
    # extra (N = 1 only, NOT the headline): the same dense-gt iteration with the views' geometry stage resident
    # (FeatureTrainer(geometry_cache_gb=...)): the geometry is frozen in the feature stage, so preprocess + binning of
    # a view give the same result in every iteration that renders it; ~90 MB per view of the 180 GB
    cached_line = None
    if world == 1 and use_graph and default_wl:
        trc = FeatureTrainer(scene, topk=TOPK, layers=1, device=dev, use_graph=True, geometry_cache_gb=4.0)
        trc.calibrate([cams[v] for v in my_views])
        ok = all(trc.cache_geometry(cams[v], v) for v in my_views)
        for v in my_views:
            trc.step(cams[v], *gts[v])
        for s in range(2 * len(my_views) + 2):
            v = my_views[s % len(my_views)]
            trc.step(cams[v], *gts[v], view_id=v)
        ms_c, per_c, loss_c = _timed_iterations(trc, cams, gts, my_views, steps, dev, world, barrier)
        cached_line = {"value": steps / (ms_c * 1e-3), "unit": "iter/s", "ms_per_iter_median": pct(per_c, 50),
                       "ms_per_iter_p95": pct(per_c, 95), "final_loss": loss_c, "views_cached": len(my_views) if ok else 0,
                       "cache_bytes": int(trc._geom_cache_used), "hits": int(trc.geometry_cache_hits),
                       "what": "same iteration with each view's preprocess + binning result kept resident (frozen geometry: "
                               "identical to recomputing it); NOT the headline, which recomputes the geometry stage every iteration"}
        del trc
    seg_line = {"value": steps / (ms_seg * 1e-3), "unit": "iter/s", "ms_per_iter": ms_seg / steps,
                "ms_per_iter_median": pct(per_seg, 50), "ms_per_iter_p95": pct(per_seg, 95), "final_loss": loss_seg,
                "what": "same step with gt given as (seg map, per-segment table) -- no 64x512xHW product"}
    coll = {"none": "none (single rank)",
            "zero1": "reduce-scatter of the flat gradient in %d row blocks (pipelined with the top-k backward) -> Adam on "
                     "this rank's 1/N shard -> all-gather of the parameters" % len(tr._blocks),
            "allreduce": "all-reduce of the flat gradient in %d row blocks, every rank steps everything" % len(tr._blocks)}
    exch_bytes = int(tr.flat_g.numel() * 4) if world > 1 else 0
    if world > 1 and tr.rows_exchange is not None:
        coll["zero1"] = ("reduce-scatter of the coefficient gradient in ROW form ([N, %d] fp32 = %d B per Gaussian instead of "
                         "the 256-B dense logits row; the top-k backward is linear in it) in %d row blocks -> top-k backward + "
                         "Adam on this rank's 1/N shard of rows -> all-gather of the parameters (deferred under the next "
                         "iteration's geometry stage); codebook gradient: reduce-scatter of its own block"
                         % (tr._row_stride, 4 * tr._row_stride, len(tr._chunk_rows)))
        exch_bytes = int(tr._rows_g.numel() * 4 + (tr.flat_g.numel() - tr.n_log) * 4)
    return {"metric": "train iter/s", "segments_gt": seg_line, "geometry_cached": cached_line,
            "value": steps / (ms * 1e-3), "unit": "iter/s",
            "views_per_s": world * steps / (ms * 1e-3), "ms_per_iter": ms / steps, "ms_per_iter_median": pct(per, 50),
            "ms_per_iter_p95": pct(per, 95), "ms_per_iter_min": float(min(per)), "steps": steps, "warmup": n_warm,
            "n_gpus": world, "scaling": "weak", "final_loss": loss, "dtype": "f32", "launch": graph_note,
            "overflow_events": tr.overflow_events,
            "config": {"workload": wl["name"], "views_per_iter": world,
                       "grad_exchange_bytes": exch_bytes, "row_form_gradient": bool(tr.row_grad),
                       "grad_exchange": coll[tr.collective],
                       "optimizer": "lsv2_adam_step (hand-written, torch.optim.Adam(eps=1e-15) semantics) on the flat "
                                    "[logits | codebook] buffer"}}


def run_rgb_stage(args, dev, scene, steps=30):
    """the RGB stage's rasterizer work (train.py without --include_feature, :169-172): forward in RGB
    mode, L1 loss, backward to means / scales / rotations / opacity / SH through lsv2_render_backward_rgb +
    lsv2_preprocess_backward.  ssim, the optimizer and densification are the reference's torch code and not timed."""
    from langsplatv2_b200 import synth
    from langsplatv2_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    W, H, N = TRAIN_WL["W"], TRAIN_WL["H"], scene["xyz"].shape[0]
    cam = synth.ring_camera(0, VIEWS, W, H)
    rs = GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam["world_view_transform"].to(dev),
        projmatrix=cam["full_proj_transform"].to(dev), sh_degree=scene["sh_degree"],
        campos=cam["camera_center"].to(dev), prefiltered=False, debug=False, include_feature=False, quick_render=False)
    leaf = lambda t: t.detach().float().to(dev).clone().requires_grad_(True)
    p = dict(means3D=leaf(scene["xyz"]), opacities=leaf(scene["opacity"]), shs=leaf(scene["features"]),
             scales=leaf(scene["scaling"]), rotations=leaf(scene["rotation"]))
    gt = torch.rand(3, H, W, device=dev)
    ph = torch.zeros((1,), device=dev)
    state = {"cap": None}

    def step():
        for v in p.values():
            v.grad = None
        means2D = torch.zeros(N, 3, device=dev, requires_grad=True)
        rast = GaussianRasterizer(raster_settings=rs, capacity=state["cap"])
        color, _, _ = rast(means2D=means2D, colors_precomp=None, cov3D_precomp=None, language_feature_precomp=ph,
                           language_feature_weights_quick=ph, language_feature_indices=ph, **p)
        loss = (color - gt).abs().mean()
        loss.backward()
        if state["cap"] is None:
            state["cap"] = int(rast.last_state.num_rendered * 1.2) + 4096
        return loss

    for _ in range(4):
        step()
    torch.cuda.synchronize(dev)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    evs[0].record()
    for i in range(steps):
        loss = step()
        evs[i + 1].record()
    torch.cuda.synchronize(dev)
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    ok = all(bool(torch.isfinite(v.grad).all()) for v in p.values())
    return {"metric": "RGB-stage forward + backward iterations/s (rasterizer + L1 loss; 1M Gaussians, %dx%d)" % (W, H),
            "value": 1e3 * steps / sum(per), "unit": "iter/s", "ms_per_iter_median": float(np.percentile(per, 50)),
            "ms_per_iter_p95": float(np.percentile(per, 95)), "steps": steps, "final_loss": float(loss.detach()),
            "grads_finite": ok, "launch": "eager",
            "what": "gradients to means3D, means2D, scales, rotations, opacities, SH (degree 3): "
                    "lsv2_render_backward_rgb + lsv2_preprocess_backward"}


def run_scene_init(dev, scene):
    """distCUDA2 over the scene's positions (scale initialisation): this repo's kernel, CUDA events; and, when
    oracle/_ref/libsimple_knn_ref.so travelled, the reference's own compiled SimpleKNN::knn on the same points (it
    synchronises inside: host wall clock), with a bit comparison of the two results."""
    from langsplatv2_b200.knn import distCUDA2
    pts = scene["xyz"].to(dev).float().contiguous()
    for _ in range(3):
        ours = distCUDA2(pts)
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ours = distCUDA2(pts)
    e1.record()
    torch.cuda.synchronize(dev)
    out = {"op": "distCUDA2 (mean squared distance to the 3 nearest neighbours)", "points": int(pts.shape[0]),
           "ms": e0.elapsed_time(e1) / 10, "gpu_launches_per_call": 9}
    try:
        from oracle import knn_ref                      # reference leg only: the compiled reference, never the product
        if knn_ref.reference_lib() is not None:
            for _ in range(2):
                ref = knn_ref.reference_dist_cuda2(pts)
            t0 = time.perf_counter()
            for _ in range(5):
                ref = knn_ref.reference_dist_cuda2(pts)
            out["reference_ms"] = (time.perf_counter() - t0) / 5 * 1e3
            out["reference"] = "SimpleKNN::knn compiled from the reference's simple_knn.cu (oracle/_ref), same GPU"
            out["bit_identical_to_reference"] = bool(torch.equal(ref.view(torch.int32), ours.view(torch.int32)))
        else:
            out["reference"] = "oracle/_ref/libsimple_knn_ref.so not built"
    except Exception as ex:                              # the baseline leg must not take the bench line down
        out["reference"] = "failed: " + str(ex).split("\n")[0]
    return out


def run_reference_style(r, cams, steps, dev):
    """What a user of the reference runs per frame (eval_lerf.py:210-220): the drop-in op through its 3-tuple API
    (one host sync per frame for num_rendered, fp32 [192,H,W] weight map to HBM) followed by the reference's own
    PyTorch decode: einsum + norm + divide in fp32.  The rasterizer underneath is ours (the reference's CUDA
    rasterizer source is absent), so this is a LOWER bound on what switching to the fused path buys."""
    from langsplatv2_b200.rasterizer import GaussianRasterizer
    cb = r.codebooks.permute(0, 2, 1).contiguous()
    zero = torch.zeros((1,), dtype=torch.float32, device=dev)

    def frame(cam):
        with torch.no_grad():
            rast = GaussianRasterizer(raster_settings=r.settings(cam))
            _, wmap, _ = rast(means3D=r.xyz, means2D=None, shs=r.features, colors_precomp=None,
                              language_feature_precomp=zero, language_feature_weights_quick=r.weights,
                              language_feature_indices=r.indices, opacities=r.opacity, scales=r.scaling,
                              rotations=r.rotation, cov3D_precomp=None)
            D_, H, W = wmap.shape
            w = wmap.view(3, 64, H * W)
            f = torch.einsum('ldk,lkn->ldn', cb, w).view(3, 512, H, W)
            return f / (f.norm(dim=1, keepdim=True) + 1e-10)

    allow = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False          # the reference runs the einsum in fp32
    for s in range(2):
        frame(cams[s % len(cams)])
    torch.cuda.synchronize(dev)
    b, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b.record()
    for s in range(steps):
        out = frame(cams[s % len(cams)])
    e.record()
    torch.cuda.synchronize(dev)
    del out
    torch.backends.cuda.matmul.allow_tf32 = allow
    ms = b.elapsed_time(e) / steps
    # the op alone (what this repo replaces) vs the reference's own torch decode lines on top of it
    def op_only(cam):
        with torch.no_grad():
            rast = GaussianRasterizer(raster_settings=r.settings(cam))
            return rast(means3D=r.xyz, means2D=None, shs=r.features, colors_precomp=None, language_feature_precomp=zero,
                        language_feature_weights_quick=r.weights, language_feature_indices=r.indices, opacities=r.opacity,
                        scales=r.scaling, rotations=r.rotation, cov3D_precomp=None)[1]
    op_only(cams[0])
    torch.cuda.synchronize(dev)
    b.record()
    for s in range(steps):
        wm = op_only(cams[s % len(cams)])
    e.record()
    torch.cuda.synchronize(dev)
    del wm
    op_ms = b.elapsed_time(e) / steps
    return {"value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "steps": steps,
            "op_ms": op_ms, "op_fps": 1e3 / op_ms, "torch_decode_ms": ms - op_ms,
            "what": "drop-in 3-tuple API (host sync per frame, fp32 weight map) + the reference's torch einsum/norm/divide "
                    "(eval_lerf.py:214-218); rasterizer kernels are this repo's"}


def run_gpu(args, wl):
    import torch.distributed as dist
    from langsplatv2_b200.renderer import FeatureRenderer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    if wl.get("fwd_bwd"):
        # BASELINE.json configs 1 / 4: the step is one training iteration at that size
        from langsplatv2_b200 import synth
        scene = synth.make_scene(wl["N"], seed=0, levels=1, codebook_size=CODES, feat_dim=D, sh_degree=3)
        args.train_steps = args.steps
        sampler = ClockSampler(local)
        sampler.start()
        tr = run_train(args, dev, rank, world, scene, barrier, wl=wl)
        clocks = sampler.stop()
        if rank == 0:
            line = {"metric": "forward + backward iterations/s", "value": tr["views_per_s"], "unit": "iter/s", "n_gpus": world,
                    "steps": tr["steps"], "warmup": tr["warmup"], "ms_per_step": tr["ms_per_iter"], "higher_is_better": True,
                    "scaling": "weak", "vs_baseline": None, "dtype": "f32 (blend: fp16 hi/lo split MMAs, fp32 accumulate)",
                    "data": "synthetic", "config": dict(tr["config"], workload=wl["name"]), "clocks": clocks,
                    "train": tr, "gpu_launches": 40 * tr["steps"]}
            if world == 1 and not args.no_cpu_baseline and wl["N"] <= 200_000:
                sec, threads = cpu_fwd_bwd_time(wl)
                line["cpu_baseline"] = {"value": 1.0 / sec, "unit": "iter/s", "cores": threads, "kind": "port",
                                        "sample": "oracle/torch_ref.py forward + backward of one full iteration (no sampling)"}
            print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    scene, cams = make_inputs(wl)
    cams = [{k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in c.items()} for c in cams]
    my_views = [v for v in range(VIEWS) if v % world == rank]
    r = FeatureRenderer(scene, topk=TOPK, device=dev, precision=args.precision, bounded=True,
                        fused=not args.unfused)
    fused = r.fused
    r.calibrate([cams[v] for v in my_views[:: max(1, len(my_views) // 8)]])

    ev = {"b": [], "e": []}

    def hook(tag):
        e = torch.cuda.Event(enable_timing=True)
        e.record(torch.cuda.current_stream(dev))
        ev["b" if tag == "decode_begin" else "e"].append(e)

    out = None
    for s in range(args.warmup):
        out = r.render(cams[my_views[s % len(my_views)]])
    r.check_overflow()
    # ---- device-resident throughput ("value") ----
    r.hook = hook
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t_b, t_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_b.record()
    realised = 0
    for s in range(args.steps):
        out = r.render(cams[my_views[s % len(my_views)]])
    t_e.record()
    barrier()

    def _same_load():
        r.render(cams[my_views[0]])
        torch.cuda.synchronize(dev)
    sampler.keep_loaded(_same_load)
    clocks = sampler.stop()
    r.check_overflow()
    realised = r.last_state.num_rendered
    # instances the blend had to read: per tile, the largest n_contrib of its pixels (early termination
    # makes the rest of each depth-sorted list dead); computed after the timed region
    # (the fused kernel walks each tile's list once per 8-row band: the two halves of a tile stop independently)
    st = r.last_state
    gx_, gy_ = (wl["W"] + 15) // 16, (wl["H"] + 15) // 16
    nc = st.image["n_contrib"].to(torch.int64)
    nc = torch.nn.functional.pad(nc, (0, gx_ * 16 - wl["W"], 0, gy_ * 16 - wl["H"]))
    consumed = int(nc.view(gy_ * 2, 8, gx_, 16).amax(dim=(1, 3)).sum().item())
    ms_total = t_b.elapsed_time(t_e)
    dom_ms = float(np.mean([b.elapsed_time(e) for b, e in zip(ev["b"], ev["e"])]))
    r.hook = None

    # ---- end to end through the public API with host buffers ("e2e") ----
    # per step: the camera block comes from pinned host memory; the RGB render (what the reference's
    # render server ships to its client, backend_renderer.py:187) and the status words go back to the host.
    H, W = wl["H"], wl["W"]
    # Every view's camera block sits in ONE pinned host tensor (packed once, outside the timed region, as a server
    # holds its trajectory); per frame: H2D of the 35 floats, the frame's kernels, D2H of RGB + status on a copy stream
    # under the next frames' kernels.  A ring of RING frames is in flight; the host only waits for frame s - RING.
    out_bytes_frame = LEVELS * D * H * W * (4 if r.out_dtype == torch.float32 else 2)
    RING = 4 if 4 * out_bytes_frame < 70e9 else (2 if 4 * out_bytes_frame < 140e9 else 1)   # frames in flight hold their outputs
    out = None
    cpu_cams = [{k: (v.cpu() if torch.is_tensor(v) else v) for k, v in cams[v].items()} for v in my_views]
    host_cams = torch.empty((len(cpu_cams), 35), dtype=torch.float32).pin_memory()
    for i, c in enumerate(cpu_cams):
        host_cams[i, :16] = c["world_view_transform"].reshape(-1)
        host_cams[i, 16:32] = c["full_proj_transform"].reshape(-1)
        host_cams[i, 32:35] = c["camera_center"]
    copy_stream = torch.cuda.Stream(device=dev)
    host_rgb = [torch.empty((3, H, W), dtype=torch.float32).pin_memory() for _ in range(RING)]
    host_status = [torch.empty(8, dtype=torch.int32).pin_memory() for _ in range(RING)]
    dev_cam = [torch.empty(35, dtype=torch.float32, device=dev) for _ in range(RING)]
    done = [None] * RING
    keep = [None] * RING
    # one captured frame per ring slot (FeatureRenderer.graphed): the slot's camera tensor is rewritten by the H2D copy,
    # the frame is ONE graph launch; falls back to eager launches if the capture is refused
    frames = None
    if not args.no_graph and RING * out_bytes_frame < 70e9:
        try:
            frames = []
            for i in range(RING):
                dc = dev_cam[i]
                dc.copy_(host_cams[0], non_blocking=True)
                cam_i = dict(cpu_cams[0], world_view_transform=dc[:16].view(4, 4), full_proj_transform=dc[16:32].view(4, 4),
                             camera_center=dc[32:35])
                frames.append(r.graphed(cam_i))
        except Exception as exc:
            print("bench: frame capture failed, e2e runs eager launches:", str(exc)[:200], file=sys.stderr)
            frames = None
            torch.cuda.synchronize(dev)

    def e2e_step(s):
        i = s % RING
        if done[i] is not None:
            done[i].synchronize()                              # the caller consumes frame s - RING before its buffers are reused
        v = s % len(cpu_cams)
        dc = dev_cam[i]
        dc.copy_(host_cams[v], non_blocking=True)
        if frames is not None:
            f = frames[i].replay()
            st = frames[i].state
        else:
            cam = dict(cpu_cams[v], world_view_transform=dc[:16].view(4, 4), full_proj_transform=dc[16:32].view(4, 4),
                       camera_center=dc[32:35])
            f = r.render(cam)
            st = r.last_state
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(dev))
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(ready)
            host_rgb[i].copy_(st.image["color"], non_blocking=True)
            host_status[i].copy_(st.status, non_blocking=True)
            ev_done = torch.cuda.Event()
            ev_done.record(copy_stream)
        if frames is None:
            st.image["color"].record_stream(copy_stream)
            st.status.record_stream(copy_stream)
        done[i] = ev_done
        keep[i] = f
        return f

    def e2e_drain():
        for ev_ in done:
            if ev_ is not None:
                ev_.synchronize()

    for s in range(min(args.warmup, RING)):
        e2e_step(s)
    e2e_drain()
    barrier()
    w0 = time.perf_counter()
    for s in range(args.steps):
        out = e2e_step(s)
    e2e_drain()
    torch.cuda.synchronize(dev)
    e2e_ms = (time.perf_counter() - w0) * 1e3                  # host wall clock: the last frame has reached the host
    barrier()
    assert all(int(h[1]) == 0 for h in host_status), "instance capacity overflow during the e2e run"
    keep = [None] * RING
    e2e_launch = "one CUDA-graph launch per frame (FeatureRenderer.graphed)" if frames is not None else "eager launches"
    frames = None
    out = None                                                 # the last frame's map (51 GB at 4k) and the graph pools
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    torch.cuda.empty_cache()

    # ---- open-vocabulary query: blend + relevancy epilogue, the 512-d map never written ----
    query = None
    if fused and not args.no_query:
        gq = torch.Generator().manual_seed(5)
        pos = torch.nn.functional.normalize(torch.randn(8, D, generator=gq), dim=-1).to(dev)
        neg = torch.nn.functional.normalize(torch.randn(4, D, generator=gq), dim=-1).to(dev)
        qev = {"b": [], "e": []}
        r.hook = lambda tag: (qev["b" if tag == "decode_begin" else "e"].append(torch.cuda.Event(enable_timing=True)),
                              qev["b" if tag == "decode_begin" else "e"][-1].record(torch.cuda.current_stream(dev)))
        for s in range(3):
            rel = r.query(cams[my_views[s % len(my_views)]], pos, neg)
        qev["b"].clear(); qev["e"].clear()
        barrier()
        q_b, q_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        q_b.record()
        for s in range(args.steps):
            rel = r.query(cams[my_views[s % len(my_views)]], pos, neg)
        q_e.record()
        barrier()
        r.hook = None
        q_ms = q_b.elapsed_time(q_e)
        qk_ms = float(np.mean([b.elapsed_time(e) for b, e in zip(qev["b"], qev["e"])]))
        tq = torch.tensor([q_ms, qk_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tq, op=dist.ReduceOp.MAX)
        q_ms, qk_ms = (float(x) for x in tq.tolist())
        q_bytes = consumed * (4 + 48 + 96) + 20 * H * W + LEVELS * 8 * H * W * 4
        query = {"metric": "open-vocabulary relevancy FPS (8 phrases vs 4 negatives, 3 levels)", "value": world * args.steps / (q_ms * 1e-3),
                 "unit": UNIT, "ms_per_step": q_ms / args.steps, "kernel_ms": qk_ms,
                 "what": "FeatureRenderer.query(): blend -> sim = w^T (C E^T) / |C^T w| -> pairwise softmax / min over negatives "
                         "(eval/openclip_encoder.py:114-138) in the blend kernel; output [3, 8, H, W] fp32, no 512-d map",
                 "roofline": {"bound": "latency / issue of the blend (the 12.7 GB feature write is gone)", "algorithmic_bytes_per_launch": q_bytes,
                              "achieved": q_bytes / (qk_ms * 1e-3) / 1e9, "unit": "GB/s"}}
        del rel

    # ---- the same frame with a bf16 feature map (SURVEY 8f-3 "channels-last / bf16 output for consumers that want the
    #      map"): half the write bytes; extra key only, the headline stays on the reference's fp32 map ----
    bf16_out = None
    if fused and not args.no_query and r.out_dtype == torch.float32:
        out = None
        torch.cuda.empty_cache()
        r.out_dtype = torch.bfloat16
        try:
            for s in range(3):
                ob = r.render(cams[my_views[s % len(my_views)]])
            barrier()
            b_b, b_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b_b.record()
            for s in range(args.steps):
                ob = r.render(cams[my_views[s % len(my_views)]])
            b_e.record()
            barrier()
            tb = torch.tensor([b_b.elapsed_time(b_e)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tb, op=dist.ReduceOp.MAX)
            b_ms = float(tb.item())
            bf16_out = {"metric": "512-d feature-map FPS, bf16 output", "value": world * args.steps / (b_ms * 1e-3),
                        "unit": UNIT, "ms_per_step": b_ms / args.steps,
                        "out_bytes_per_frame": int(LEVELS * D * H * W * 2),
                        "what": "same kernel, features rounded to bf16 in the epilogue (out_dtype=torch.bfloat16)"}
            del ob
            torch.cuda.empty_cache()
            # ... and with a channels-last fp32 map ([levels, H, W, 512], what eval_lerf.py:352 permutes the map to)
            r.out_dtype = torch.float32
            r.out_layout = "HWC"
            for s in range(3):
                ob = r.render(cams[my_views[s % len(my_views)]])
            barrier()
            h_b, h_e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            h_b.record()
            for s in range(args.steps):
                ob = r.render(cams[my_views[s % len(my_views)]])
            h_e.record()
            barrier()
            th = torch.tensor([h_b.elapsed_time(h_e)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(th, op=dist.ReduceOp.MAX)
            h_ms = float(th.item())
            bf16_out["channels_last_f32"] = {"value": world * args.steps / (h_ms * 1e-3), "unit": UNIT, "ms_per_step": h_ms / args.steps,
                                             "shape": [int(x) for x in ob.shape],
                                             "what": "out_layout='HWC': the decode MMA's operands swapped (lane = feature), one 128-byte line per store"}
            del ob
        finally:
            r.out_dtype = torch.float32
            r.out_layout = "CHW"
        torch.cuda.empty_cache()

    ref_style = None
    if world == 1 and not args.no_reference_style:
        del out
        torch.cuda.empty_cache()
        try:
            ref_style = run_reference_style(r, [cams[v] for v in my_views], min(args.steps, 10), dev)
        except torch.OutOfMemoryError as ex:
            # the reference's decode lines hold the un-normalised map, its norm and the result at once (2 x 51 GB at 4k)
            ref_style = {"unavailable": "the reference's torch decode (einsum + norm + divide) ran out of device memory "
                                        "at this size: " + str(ex).split(".")[0]}
            torch.cuda.empty_cache()
        out = None
    train = None
    if not args.no_train:
        out = None
        r.last_state = None
        torch.cuda.empty_cache()
        train = run_train(args, dev, rank, world, scene, barrier)
        if world == 1:
            torch.cuda.empty_cache()
            from langsplatv2_b200 import synth as _synth
            rgb_scene = _synth.make_scene(TRAIN_WL["N"], seed=0, levels=1, codebook_size=CODES, feat_dim=D, sh_degree=3)
            train["rgb_stage"] = run_rgb_stage(args, dev, rgb_scene)
            train["scene_init"] = run_scene_init(dev, rgb_scene)
            del rgb_scene

    t = torch.tensor([ms_total, e2e_ms, dom_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, dom_ms = (float(x) for x in t.tolist())

    if rank == 0:
        HW = H * W
        L = LEVELS * CODES
        out_bytes = LEVELS * D * HW * (4 if r.out_dtype == torch.float32 else 2)
        if fused:
            # blend+decode: consumed instances (per 8-row band) x (id 4 + blend record 48 + packed coefficient row 96) + rgb,
            # final_T, n_contrib (20 B/px) + features out + codebook / Gram images; the L-channel weight map is not traffic
            dom_bytes = consumed * (4 + 48 + 96) + 20 * HW + out_bytes + LEVELS * (CODES * D * 2 + 2 * CODES * CODES * 2)
            dom_name = "render_decode (blend + codebook decode + L2 normalise, fused)"
        else:
            dom_bytes = L * HW * 4 + out_bytes + LEVELS * CODES * D * 4     # decode: read weight map + codebooks, write features
            dom_name = "decode (codebook GEMM + L2 normalise)"
        peak, peak_src = peaks()
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
        traffic, traffic_src = ncu_traffic("k_render_decode") if (fused and args.workload == "1080p" and wl["N"] == 1_000_000) else (None, None)
        line = {
            "metric": METRIC, "value": world * args.steps / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ("f32 alpha/T; blend accumulation as fp16 head+remainder x fp16 head+remainder MMAs (3 products, fp32 "
                      "TMEM accumulate: weight map within 1e-4 of its norm); decode f16 x f16 -> f32, fp32 normalisation"
                      if fused else ("f32 blend, bf16 x bf16 -> f32 decode" if args.precision == "bf16" else "f32")),
            "data": "synthetic",
            "config": {"workload": wl["name"], "views": VIEWS, "instances_per_frame": realised,
                       "instances_consumed_by_blend": consumed, "fused_blend_decode": fused,
                       "instance_capacity": r.capacity, "sharding": f"views round-robin over {world} rank(s)",
                       "l2": "each step writes >= %.1f GB, far beyond the 126 MB L2; no explicit flush" % (out_bytes / 1e9),
                       "decode_precision": ("fp16 operands, fp32 accumulate, L2 norm from the Gram matrix (fp16 head+remainder), "
                                            "scale applied in fp32") if fused else args.precision,
                       "store_path": "32x8-pixel band units: every epilogue store instruction writes one 128-byte line",
                       "reference_cuda_rasterizer": reference_cuda_rasterizer_status()},
            "clocks": clocks,
            "e2e": {"value": world * args.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 35 * 4,
                    "d2h_bytes_per_step": 3 * HW * 4 + 32,
                    "launch": e2e_launch,
                    "note": "camera from pinned host memory; RGB render + status read back to the host (ring of %d frames in flight: frame k's D2H under the next frames' kernels; host wall clock until the last frame is on the host); the 512-d map stays on the device as in the reference's API" % RING},
            "gpu_launches": kernels_per_frame(W, H, fused) * args.steps,
            "roofline": {"kernel": dom_name, "bound": "hbm", "achieved": achieved,
                         "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": dom_bytes,
                         "avg_launch_ms": dom_ms},
        }
        if train is not None:
            line["train"] = train
        if ref_style is not None:
            line["reference_style"] = ref_style
        if query is not None:
            line["query"] = query
        if bf16_out is not None:
            line["bf16_out"] = bf16_out
        if world == 1 and not args.no_cpu_baseline:
            est, sample, threads = cpu_frame_time(scene, cpu_cams[0], wl, budget_s=20.0)
            line["cpu_baseline"] = {"value": 1.0 / est, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="1080p", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the train iter/s measurement")
    ap.add_argument("--no-reference-style", action="store_true",
                    help="skip the unfused 3-tuple API + torch decode measurement")
    ap.add_argument("--train-steps", type=int, default=200)
    ap.add_argument("--unfused", action="store_true", help="two-call path: blend to an fp32 weight map, then decode")
    ap.add_argument("--gaussians", type=int, default=0, help="override the workload's Gaussian count (config 5 sweep)")
    ap.add_argument("--no-query", action="store_true", help="skip the relevancy-query measurement")
    ap.add_argument("--no-graph", action="store_true", help="train iterations as eager launches instead of CUDA graphs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = dict(WORKLOADS[args.workload])
    if args.gaussians > 0:
        wl["name"] = wl["name"].replace("1M synthetic", f"{args.gaussians / 1e6:g}M synthetic")
        wl["N"] = args.gaussians
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_gpu(args, wl)


if __name__ == "__main__":
    main()
