import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from langsplatv2_b200.loss import _decode_bwd
D, HW = 512, 6400
g = torch.Generator(device="cuda").manual_seed(2)
C = torch.randn((1, 64, D), generator=g, device="cuda")
gt = torch.randn((D, HW), generator=g, device="cuda")
h = torch.full((64, HW), float("nan"), device="cuda")
_decode_bwd(1, 64, D, HW, None, C, gt, h, None, gt.device)
ref = (C[0].double() @ gt.double()).float()
print("h[:3,:6]\n", h[:3, :6].cpu().numpy(), "\nref\n", ref[:3, :6].cpu().numpy())
print("abs max h", float(h.abs().max()), "ref", float(ref.abs().max()), "nonzero frac", float((h != 0).float().mean()))
# simple structured inputs: C = delta on (k, d=k), gt[d][p] = d + p/1000
C2 = torch.zeros((1, 64, D), device="cuda"); C2[0, torch.arange(64), torch.arange(64) * 3 + 1] = 1.0
gt2 = (torch.arange(D, device="cuda").float()[:, None] + torch.arange(HW, device="cuda").float()[None, :] / 8192.0).contiguous()
_decode_bwd(1, 64, D, HW, None, C2, gt2, h, None, gt.device)
print("delta test h[:4,:5]\n", h[:4, :5].cpu().numpy(), "\nexpect rows d=1,4,7,10: ", gt2[[1, 4, 7, 10], :5].cpu().numpy())
print("h[:4,126:131]\n", h[:4, 126:131].cpu().numpy())
