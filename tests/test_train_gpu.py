"""Feature-stage training step (train.py --include_feature) on the CUDA path vs the fp64 PyTorch restatement.

The checker here is oracle/torch_ref.py end to end: softmax/top-k/renorm -> preprocess -> binning -> blend ->
decode one layer -> normalise -> cosine loss, differentiated by torch autograd in float64.  Stated tolerance for the
gradients: max-abs <= 2e-3 x max|ref| (fp32 blend + fp32 atomics in the coefficient backward; a pixel whose
alpha/T threshold flips between fp64 exp and the fp32 polynomial moves one contribution).
"""
import numpy as np
import pytest
import torch

from oracle import torch_ref
from langsplatv2_b200 import softmax_to_topk_soft_code
from langsplatv2_b200.trainer import FeatureTrainer, cos_loss, synthetic_gt
from tests import helpers as hp

pytestmark = pytest.mark.gpu


def test_topk_soft_code_backward_matches_autograd():
    rng = np.random.default_rng(3)
    logits = rng.standard_normal((5000, 64)).astype(np.float32) * 2.0
    g = rng.standard_normal((5000, 64)).astype(np.float32)
    x64 = torch.tensor(logits, dtype=torch.float64, requires_grad=True)
    (torch_ref.topk_soft_code(x64, 4) * torch.tensor(g, dtype=torch.float64)).sum().backward()
    x = torch.tensor(logits, device="cuda", requires_grad=True)
    y = softmax_to_topk_soft_code(x, 4)
    (y * torch.tensor(g, device="cuda")).sum().backward()
    ref = x64.grad.numpy()
    got = x.grad.cpu().numpy()
    assert np.max(np.abs(got - ref)) <= 2e-6 * max(1.0, np.abs(ref).max())
    # rows sum to ~0 (softmax Jacobian) and the non-selected channels only see the tiny coupling term
    assert np.abs(got.sum(1)).max() < 1e-5


@pytest.mark.parametrize("normalize", [True, False])
def test_fused_cos_loss_matches_torch_autograd(normalize):
    from langsplatv2_b200.loss import normalized_cos_loss
    g = torch.Generator(device="cpu").manual_seed(5)
    D, H, W = 512, 37, 53
    feat = torch.randn((D, H, W), generator=g) * 3.0
    gt = torch.randn((D, H, W), generator=g)
    gt = gt / gt.norm(dim=0, keepdim=True)
    mask = (torch.rand((1, H, W), generator=g) > 0.3).float()
    feat[:, 0, 0] = 0.0                                    # a dead pixel: 0 / (0 + 1e-10)
    f64 = feat.double().requires_grad_(True)
    x = f64 / (f64.norm(dim=0, keepdim=True) + 1e-10) if normalize else f64
    ref = cos_loss(x * mask.double(), gt.double() * mask.double())
    ref.backward()
    fc = feat.cuda().requires_grad_(True)
    out = normalized_cos_loss(fc, gt.cuda(), mask.cuda(), normalize=normalize)
    (out * 2.0).backward()                                 # a non-unit upstream gradient
    assert abs(float(out) - float(ref)) <= 2e-6
    got = fc.grad.cpu().numpy() / 2.0
    want = f64.grad.numpy()
    assert np.max(np.abs(got - want)) <= 1e-5 * np.abs(want).max() + 1e-12


@pytest.mark.parametrize("normalize", [True, False])
def test_gram_decode_cos_loss_matches_torch_autograd(normalize):
    """decode + normalise + cos loss without the feature map (lsv2_gram_loss_*) vs fp64 autograd of the reference's
    expressions (scene/gaussian_model.py:538, train.py:157-164)."""
    from langsplatv2_b200.loss import decode_cos_loss
    g = torch.Generator(device="cpu").manual_seed(8)
    H, W = 29, 41
    wmap = torch.rand((64, H, W), generator=g) * (torch.rand((64, H, W), generator=g) > 0.6)
    wmap[:, 0, 0] = 0.0                                     # dead pixel
    cb = torch.randn((1, 64, 512), generator=g)
    gt = torch.randn((512, H, W), generator=g)
    gt = gt / gt.norm(dim=0, keepdim=True)
    mask = (torch.rand((1, H, W), generator=g) > 0.3).float()
    w64 = wmap.double().requires_grad_(True)
    c64 = cb.double().requires_grad_(True)
    feat = (c64[0].T @ w64.view(64, -1)).view(512, H, W)
    if normalize:
        feat = feat / (feat.norm(dim=0, keepdim=True) + 1e-10)
    ref = cos_loss(feat * mask.double(), gt.double() * mask.double())
    ref.backward()
    wc = wmap.cuda().requires_grad_(True)
    cc = cb.cuda().requires_grad_(True)
    out = decode_cos_loss(wc, cc, gt.cuda(), mask.cuda(), normalize=normalize)
    (out * 3.0).backward()
    assert abs(float(out.detach()) - float(ref.detach())) <= 5e-6
    live = torch.ones((H, W), dtype=torch.bool)
    live[0, 0] = False                                      # the dead pixel's gradient is ~1e18 on both sides
    gw, rw = wc.grad.cpu().numpy()[:, live.numpy()] / 3.0, w64.grad.numpy()[:, live.numpy()]
    assert np.max(np.abs(gw - rw)) <= 2e-5 * np.abs(rw).max()
    gc, rc = cc.grad.cpu().numpy() / 3.0, c64.grad.numpy()
    assert np.max(np.abs(gc - rc)) <= 2e-5 * np.abs(rc).max()


def _reference_grads(scene, cam, gt, mask, topk):
    t64 = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float64)
    osc, view = hp.oracle_scene(scene), hp.oracle_view(cam)
    W, H = view["W"], view["H"]
    pre = torch_ref.preprocess(t64(osc["means3d"]), t64(osc["opacities"]), t64(osc["scales"]), t64(osc["rotations"]),
                               t64(view["viewmatrix"]), t64(view["projmatrix"]), t64(view["campos"]), W, H,
                               view["tanfovx"], view["tanfovy"], shs=t64(osc["shs"]), sh_degree=3)
    binned = torch_ref.bin_instances(pre, W, H)
    logits = scene["logits"][:, :64].double().clone().requires_grad_(True)
    cb = scene["codebooks"][:1].double().clone().requires_grad_(True)
    dense = torch_ref.topk_soft_code(logits, topk)
    out = torch_ref.render(pre, binned, W, H, torch.zeros(3), coeff=dense)
    feat = torch_ref.decode(out["wmap"], cb, normalize=False)[0]
    feat = feat / (feat.norm(dim=0, keepdim=True) + 1e-10)
    loss = cos_loss(feat * mask.double().cpu(), gt.double().cpu() * mask.double().cpu())
    loss.backward()
    return float(loss), logits.grad.numpy(), cb.grad.numpy()


@pytest.mark.parametrize("fused_loss,gram_loss", [(True, True), (True, False), (False, False)])
def test_train_step_gradients_match_fp64_autograd(fused_loss, gram_loss):
    N, W, H = 3000, 96, 64
    scene, cam = hp.make_case(N, W, H, seed=11, levels=1, scale_boost=2.5)
    gt, mask = synthetic_gt(H, W, device="cuda")
    ref_loss, ref_dl, ref_dc = _reference_grads(scene, cam, gt, mask, 4)
    tr = FeatureTrainer(scene, topk=4, layers=1, fused_loss=fused_loss, gram_loss=gram_loss)
    loss = tr.loss(cam, gt, mask)
    loss.backward()
    assert abs(float(loss) - ref_loss) <= 1e-4
    dl = tr.logits.grad.cpu().numpy()
    dc = tr.codebooks.grad.cpu().numpy()
    assert np.abs(ref_dl).max() > 0 and np.abs(ref_dc).max() > 0
    assert np.max(np.abs(dl - ref_dl)) <= 2e-3 * np.abs(ref_dl).max()
    assert np.max(np.abs(dc - ref_dc)) <= 2e-3 * np.abs(ref_dc).max()


def test_train_steps_reduce_the_loss():
    N, W, H = 20000, 160, 96
    scene, cam = hp.make_case(N, W, H, seed=12, levels=1, scale_boost=2.5)
    gt, mask = synthetic_gt(H, W, device="cuda")
    tr = FeatureTrainer(scene, topk=4, layers=1, lr=0.05)
    losses = [float(tr.step(cam, gt, mask)) for _ in range(15)]
    assert all(np.isfinite(losses))
    assert losses[-1] < losses[0] - 0.05, losses


@pytest.mark.parametrize("D,H,W", [(512, 64, 100), (64, 64, 100), (512, 300, 404), (200, 37, 44)])
def test_codebook_gradient_tensor_core_3xtf32_matches_fp64(D, H, W):
    """dC[k][d] = sum_p w[k][p] dF[d][p] on the tcgen05 path (csrc/gemm_tf32x3.cu: fp32 operands split hi + lo, three
    tf32 products, fp32 accumulate) against a float64 product.  Stated tolerance: 2e-6 x sum_p |w||dF| per element
    (the dropped lo*lo term is <= 2^-20 per product), i.e. fp32-GEMM grade, as the reference's fp32 einsum backward
    (scene/gaussian_model.py:538).  HW % 4 == 0 on every case here so the tensor-core path is the one that runs; the
    last case has a ragged row count (200 of 256) and a ragged last 32-pixel step."""
    from langsplatv2_b200.loss import _decode_bwd
    g = torch.Generator(device="cuda").manual_seed(D + H)
    HW = H * W
    assert HW % 4 == 0
    w = torch.randn((64, HW), generator=g, device="cuda") * torch.exp(2.0 * torch.randn((64, 1), generator=g, device="cuda"))
    w = w * (torch.rand((64, HW), generator=g, device="cuda") > 0.5)
    dF = torch.randn((D, HW), generator=g, device="cuda") * torch.exp(torch.randn((D, 1), generator=g, device="cuda"))
    dC = torch.zeros((1, 64, D), dtype=torch.float32, device="cuda")
    _decode_bwd(1, 64, D, HW, w, None, dF, None, dC, w.device)
    ref = w.double() @ dF.double().T
    bound = w.double().abs() @ dF.double().abs().T
    err = (dC[0].double() - ref).abs()
    assert bool((err <= 2e-6 * bound + 1e-30).all()), float((err / (bound + 1e-30)).max())
    assert float(err.max()) <= 1e-5 * float(ref.abs().max())


@pytest.mark.parametrize("normalize", [True, False])
def test_segment_ground_truth_loss_equals_dense(normalize):
    """loss.decode_cos_loss_segments (gt = table[seg], unlabelled pixels seg = -1; scene/cameras.py:59-96) against fp64
    autograd of the reference's dense expressions -- same tolerance as the dense Gram path."""
    from langsplatv2_b200.loss import decode_cos_loss_segments
    g = torch.Generator(device="cpu").manual_seed(21)
    H, W, S = 40, 52, 23
    wmap = torch.rand((64, H, W), generator=g) * (torch.rand((64, H, W), generator=g) > 0.6)
    cb = torch.randn((1, 64, 512), generator=g)
    table = torch.randn((S, 512), generator=g)
    table = table / table.norm(dim=1, keepdim=True)
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    seg = ((xs // 9 + 5 * (ys // 7)) % S).to(torch.int32)
    seg[torch.rand((H, W), generator=g) < 0.1] = -1                     # unlabelled pixels
    seg[3, 5:40] = torch.randint(0, S, (35,), generator=g).to(torch.int32)   # a row with mixed segments inside warps
    valid = seg >= 0
    gt = table[seg.clamp(min=0).long()].permute(2, 0, 1) * valid
    mask = valid.unsqueeze(0).double()
    w64 = wmap.double().requires_grad_(True)
    c64 = cb.double().requires_grad_(True)
    feat = (c64[0].T @ w64.view(64, -1)).view(512, H, W)
    if normalize:
        feat = feat / (feat.norm(dim=0, keepdim=True) + 1e-10)
    ref = cos_loss(feat * mask, gt.double() * mask)
    ref.backward()
    wc = wmap.cuda().requires_grad_(True)
    cc = cb.cuda().requires_grad_(True)
    out = decode_cos_loss_segments(wc, cc, seg.cuda(), table.cuda(), normalize=normalize)
    (out * 2.0).backward()
    assert abs(float(out.detach()) - float(ref.detach())) <= 5e-6
    live = (wmap.abs().sum(0) > 0).numpy()
    gw, rw = wc.grad.cpu().numpy()[:, live] / 2.0, w64.grad.numpy()[:, live]
    assert np.max(np.abs(gw - rw)) <= 2e-5 * np.abs(rw).max()
    gc, rc = cc.grad.cpu().numpy() / 2.0, c64.grad.numpy()
    assert np.max(np.abs(gc - rc)) <= 2e-5 * np.abs(rc).max()
