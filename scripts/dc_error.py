"""Error of the codebook-gradient product vs float64, tensor-core (default) or CUDA-core (LSV2_DC_SIMT=1) path."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from langsplatv2_b200.loss import _decode_bwd
for D, HW in [(512, 6400), (512, 121200), (512, 762048), (64, 762048)]:
    g = torch.Generator(device="cuda").manual_seed(1)
    w = torch.randn((64, HW), generator=g, device="cuda") * (torch.rand((64, HW), generator=g, device="cuda") > 0.5)
    dF = torch.randn((D, HW), generator=g, device="cuda")
    dC = torch.zeros((1, 64, D), device="cuda")
    _decode_bwd(1, 64, D, HW, w, None, dF, None, dC, w.device)
    ref = w.double() @ dF.double().T
    bound = w.double().abs() @ dF.double().abs().T
    err = (dC[0].double() - ref).abs()
    allow = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    t = (w @ dF.T).double()
    print(D, HW, "max err / max|ref| %.3e" % float(err.max() / ref.abs().max()), "max err/bound %.3e" % float((err / bound).max()),
          "| torch fp32 matmul: %.3e" % float((t - ref).abs().max() / ref.abs().max()))
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(5):
        _decode_bwd(1, 64, D, HW, w, None, dF, None, dC, w.device)
    t1.record(); torch.cuda.synchronize()
    print("   ms per call %.3f" % (t0.elapsed_time(t1) / 5))

print("h = C gt (dW product)")
for D, HW in [(512, 6400), (512, 762048), (512, 1189 * 4)]:
    g = torch.Generator(device="cuda").manual_seed(2)
    C = torch.randn((1, 64, D), generator=g, device="cuda")
    gt = torch.randn((D, HW), generator=g, device="cuda")
    h = torch.full((64, HW), float("nan"), device="cuda")
    _decode_bwd(1, 64, D, HW, None, C, gt, h, None, gt.device)
    ref = C[0].double() @ gt.double()
    err = (h.double() - ref).abs()
    print(D, HW, "max err / max|ref| %.3e" % float(err.max() / ref.abs().max()), "nan", int(torch.isnan(h).sum()))
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(5):
        _decode_bwd(1, 64, D, HW, None, C, gt, h, None, gt.device)
    t1.record(); torch.cuda.synchronize()
    print("   ms per call %.3f" % (t0.elapsed_time(t1) / 5))
