"""``ldmb200_groupnorm_static`` — the allocation-free GroupNorm entry point (statistics workspace in a file-scope device
array inside the library; include/ldmb200.h).  Collected last on purpose: it compares the
new entry point with ``ldmb200_groupnorm`` through the C-ABI, on two calls in a row, and passes alone in a fresh process
as well as after the rest of the suite."""
import os
import sys

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(__file__))
from dsml_thesis_b200 import _cabi  # noqa: E402
from dsml_thesis_b200._cabi import BF16, F16, F32  # noqa: E402

import kernel_harness as K  # noqa: E402

OUT_TOL = {F32: 1e-5, BF16: 4e-3, F16: 5e-4}   # the bounds tests/test_kernels_gpu.py uses for these output types
MAX_BATCH = 64                                  # LDMB200_GN_STATIC_MAX_BATCH


def _static(x1, x2, gamma, beta, B, hw, eps, silu, out_dtype, want_raw):
    c = x1.shape[1] + (x2.shape[1] if x2 is not None else 0)
    tdt = K.TORCH_DT[out_dtype]
    y = torch.full((B * hw, c), float("nan"), dtype=tdt, device=K.DEV)
    raw = torch.full((B * hw, c), float("nan"), dtype=tdt, device=K.DEV) if want_raw else None
    _cabi.call("ldmb200_groupnorm_static", x1.data_ptr(), x1.shape[1], K.ptr(x2), x2.shape[1] if x2 is not None else 0,
               gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), K.ptr(raw), out_dtype, B, hw, 32, float(eps), int(silu),
               K.stream())
    torch.cuda.synchronize()
    return y, raw


@pytest.mark.parametrize("B,H,c1,c2,silu,out_dtype", [
    (2, 16, 160, 0, 1, F16),       # ResBlock.in_layers at the 160-channel level
    (3, 8, 640, 320, 1, BF16),     # virtual concat whose groups straddle the two tensors, raw copy wanted
    (64, 8, 64, 0, 0, F32),        # the largest batch the built-in workspace holds; Normalize (no SiLU)
    (1, 32, 320, 0, 1, F16),       # many row splits for one sample
])
def test_groupnorm_static_matches_groupnorm_on_two_calls_in_a_row(B, H, c1, c2, silu, out_dtype):
    hw = H * H
    g = torch.Generator(device="cpu").manual_seed(1234 + B + c1)
    mk = lambda *s: torch.randn(*s, generator=g).to(K.DEV)
    c = c1 + c2
    gamma, beta = 1 + 0.1 * mk(c), 0.1 * mk(c)
    want_raw = c2 > 0
    for call in range(2):   # second call: new inputs over the previous call's leftover workspace contents
        x1 = mk(B * hw, c1) * (1.0 + call) + 0.5 * call
        x2 = mk(B * hw, c2) - 0.3 if c2 else None
        ref_y, ref_raw = K.groupnorm(x1, x2, gamma, beta, B, hw, 1e-5, silu, out_dtype, want_raw)
        y, raw = _static(x1, x2, gamma, beta, B, hw, 1e-5, silu, out_dtype, want_raw)
        assert torch.isfinite(y.float()).all()
        assert K.rel_l2(y, ref_y) <= OUT_TOL[out_dtype]
        assert torch.equal(y, ref_y)                       # same kernels, same fixed-order reduction
        if want_raw:
            assert torch.equal(raw, ref_raw)
        xc = torch.cat([x1, x2], 1) if x2 is not None else x1
        t = F.group_norm(K.from_cl(xc, B, H, H).double(), 32, gamma.double(), beta.double(), 1e-5)
        t = K.to_cl(F.silu(t) if silu else t)
        assert K.rel_l2(y, t) <= OUT_TOL[out_dtype]


def test_groupnorm_static_rejects_a_batch_beyond_its_workspace():
    B, hw, c = MAX_BATCH + 1, 16, 32
    x = torch.zeros(B * hw, c, device=K.DEV)
    gb = torch.ones(c, device=K.DEV)
    y = torch.empty(B * hw, c, device=K.DEV)
    with pytest.raises(_cabi.Ldmb200Error, match="LDMB200_GN_STATIC_MAX_BATCH"):
        _cabi.call("ldmb200_groupnorm_static", x.data_ptr(), c, None, 0, gb.data_ptr(), gb.data_ptr(), y.data_ptr(), None,
                   F32, B, hw, 32, 1e-5, 0, K.stream())
