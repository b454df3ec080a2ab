"""lsv2_decode_static: the tensor-core codebook decode with its scratch in a static device array inside liblsv2.so
(no caller scratch, no allocation) -- for callers that cannot call cudaMalloc.  No counterpart in the reference.
Checked through the C-ABI against the existing entry point lsv2_decode
(precision = LSV2_DECODE_BF16, caller-provided scratch) on two calls in a row with different codebooks, and against
the CPU checker within the tolerance the other decode tests use (4e-3 on the unit-norm feature).

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle
from langsplatv2_b200 import _abi

pytestmark = pytest.mark.gpu
TOL = 4e-3          # tests/test_parity_gpu.py::test_decode_bf16_tensor_core
MAX_LEVELS = 3      # LSV2_DECODE_STATIC_MAX_LEVELS


def _decode_existing(L_, levels, HW, wmap, cb, normalize, out_dtype):
    out = torch.empty((levels, 512, HW), dtype=torch.float32 if out_dtype == 0 else torch.bfloat16, device="cuda")
    nbytes = L_.lsv2_decode_scratch_bytes(levels, 64, 512)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _abi.check(L_.lsv2_decode(levels, 64, 512, HW, wmap.data_ptr(), cb.data_ptr(), 1 if normalize else 0,
                              _abi.DECODE_BF16, out_dtype, out.data_ptr(), scratch.data_ptr(), nbytes,
                              _abi.stream_ptr()), "lsv2_decode")
    return out


def _decode_static(L_, levels, HW, wmap, cb, normalize, out_dtype, out):
    return L_.lsv2_decode_static(levels, 64, 512, HW, wmap.data_ptr(), cb.data_ptr(), 1 if normalize else 0,
                                 out_dtype, out.data_ptr(), _abi.stream_ptr())


def test_decode_static_matches_lsv2_decode_on_two_calls_in_a_row():
    L_ = _abi.lib()
    g = torch.Generator(device="cuda").manual_seed(21)
    cases = [(3, 16 * 333 + 5, True, 0), (2, 4096, True, 0), (3, 1000, False, 0), (1, 128 * 7, True, 1)]
    outs = [torch.empty((lv, 512, HW), dtype=torch.float32 if od == 0 else torch.bfloat16, device="cuda")
            for lv, HW, _, od in cases]                     # every buffer exists before the first no-alloc call
    for (levels, HW, normalize, od), out in zip(cases, outs):
        # sparse non-negative weights as the blend leaves them; a fresh codebook per call: the static array must not
        # leak the previous call's image
        wmap = (torch.rand((levels * 64, HW), generator=g, device="cuda") *
                (torch.rand((levels * 64, HW), generator=g, device="cuda") < 0.1)).contiguous()
        cb = torch.randn((levels, 64, 512), generator=g, device="cuda").contiguous()
        want = _decode_existing(L_, levels, HW, wmap, cb, normalize, od)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        out.fill_(float("nan"))
        rc = _decode_static(L_, levels, HW, wmap, cb, normalize, od, out)
        torch.cuda.synchronize()
        _abi.check(rc, "lsv2_decode_static")
        assert torch.cuda.mem_get_info()[0] >= free0 - (1 << 20), "device memory shrank across a no-alloc call"
        a, b = out.float(), want.float()
        assert torch.isfinite(a).all()
        scale = 1.0 if normalize else float(b.abs().max())
        assert float((a - b).abs().max()) <= TOL * scale
        assert torch.equal(out, want), "same kernel, same inputs: expected bit-identical output"
        if normalize and od == 0:
            ref = oracle.decode(wmap.cpu().numpy(), cb.cpu().numpy(), normalize=True)
            assert float(np.max(np.abs(a.cpu().numpy() - ref.reshape(a.shape)))) <= TOL


def test_decode_static_rejects_what_its_scratch_cannot_hold():
    L_ = _abi.lib()
    levels, HW = MAX_LEVELS + 1, 256
    wmap = torch.zeros((levels * 64, HW), device="cuda")
    cb = torch.zeros((levels, 64, 512), device="cuda")
    out = torch.zeros((levels, 512, HW), device="cuda")
    rc = _decode_static(L_, levels, HW, wmap, cb, True, 0, out)
    assert rc == 3, rc                                                     # LSV2_ENOSPC
    assert b"lsv2_decode_static" in L_.lsv2_strerror(rc)
    with pytest.raises(_abi.Lsv2Error):
        _abi.check(rc, "lsv2_decode_static")
    # other codebook shapes: unsupported, not silently wrong
    assert L_.lsv2_decode_static(1, 32, 512, HW, wmap.data_ptr(), cb.data_ptr(), 1, 0, out.data_ptr(),
                                 _abi.stream_ptr()) == 4                   # LSV2_EUNSUPPORTED
    assert L_.lsv2_decode_static(1, 64, 512, HW, None, cb.data_ptr(), 1, 0, out.data_ptr(), _abi.stream_ptr()) == 1
