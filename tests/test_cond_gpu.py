"""GPU parity of the talking-face audio conditioning encoder: the fused
``ldmb200_conv1d_temporal_attention`` launch behind ``dsml_thesis_b200.Conv1DTemporalAttention`` against the committed
output of the REAL reference module (``tests/golden/cond.npz``) and the CPU oracle.  fp32 throughout: rel-L2 ≤ 1e-5."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import dsml_thesis_b200 as P  # noqa: E402
from oracle import cond_oracle as CO  # noqa: E402
from oracle import ldm_oracle as O  # noqa: E402
from conftest import GOLDEN, record_parity  # noqa: E402

HIDDEN = dict(seq_len=9, subspace_dim=64, subspace2hidden=True, hidden_dim=48)


def _model(cfg):
    sd = CO.init_cond_state_dict(cfg, 0)
    m = P.Conv1DTemporalAttention(**cfg)
    m.load_state_dict(sd, strict=True)
    return sd, m.cuda().eval()


@pytest.mark.parametrize("name,cfg", [("tf_audio", CO.TF_AUDIO), ("hidden", HIDDEN)])
def test_forward_vs_reference_golden(name, cfg):
    g = np.load(os.path.join(GOLDEN, "cond.npz"))
    _, m = _model(cfg)
    with torch.no_grad():
        y = m(torch.from_numpy(g[f"{name}/x"]).cuda())
    err = O.rel_l2(y.cpu(), torch.from_numpy(g[f"{name}/y"]))
    record_parity(f"cond/{name}/rel_l2", err)
    assert y.shape == g[f"{name}/y"].shape and err <= 1e-5, err


@pytest.mark.parametrize("B,L,D", [(1, 17, 768), (32, 17, 768), (5, 1, 32), (2, 32, 100), (3, 11, 29)])
def test_forward_vs_oracle_shapes(B, L, D):
    cfg = dict(seq_len=L, subspace_dim=D, subspace2hidden=False)
    sd, m = _model(cfg)
    x = torch.randn(B, L, D, generator=torch.Generator().manual_seed(B * 100 + L))
    ref = CO.conv1d_temporal_attention(sd, cfg, x)
    with torch.no_grad():
        y = m(x.cuda())
    assert O.rel_l2(y.cpu(), ref) <= 1e-5


def test_inference_only_and_no_cpu_fallback():
    _, m = _model(CO.TF_AUDIO)
    x = torch.randn(1, 17, 768)
    with pytest.raises(NotImplementedError):
        m(x.cuda())                          # grad mode on, trainable parameters: not this path
    with torch.no_grad(), pytest.raises(P.Ldmb200Error):
        m(x)
    with pytest.raises(NotImplementedError):
        P.Conv1DTemporalAttention(seq_len=64)


def test_install_wrapper_uses_the_live_modules_parameters():
    """The class-level wrapper install() puts on the reference class: fused launch under no_grad, the reference's own
    forward when a gradient is needed; parameters are read live (no copy to go stale)."""
    from dsml_thesis_b200.cond import accelerated_forward
    sd, donor = _model(CO.TF_AUDIO)
    calls = []

    class Ref(torch.nn.Module):               # attribute layout of modules.py:77-100
        def __init__(self):
            super().__init__()
            self.seq_len, self.subspace_dim, self.subspace2hidden = 17, 768, False
            self.attentionConvNet, self.attentionNet = donor.attentionConvNet, donor.attentionNet

        def forward(self, x):
            calls.append(1)
            return x[:, :1]

    Ref.forward = accelerated_forward(Ref.forward)
    m = Ref().cuda()
    x = torch.randn(2, 17, 768, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        y = m(x.cuda())
    assert not calls and O.rel_l2(y.cpu(), CO.conv1d_temporal_attention(sd, CO.TF_AUDIO, x)) <= 1e-5
    m(x.cuda())                                # grad enabled + trainable parameters -> reference forward
    assert calls == [1]
    with torch.no_grad():
        m.attentionNet[0].bias.add_(1.0)       # softmax is shift-invariant ...
        m.attentionNet[0].bias[3] += 2.0       # ... but not to a single-entry change
        sd2 = {k: v.detach().cpu() for k, v in m.state_dict().items()}
        y2 = m(x.cuda())
    assert O.rel_l2(y2.cpu(), CO.conv1d_temporal_attention(sd2, CO.TF_AUDIO, x)) <= 1e-5
    assert O.rel_l2(y2.cpu(), y.cpu()) > 1e-3
