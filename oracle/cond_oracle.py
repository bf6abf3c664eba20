"""CPU oracle of the talking-face audio conditioning encoder — TEST INFRASTRUCTURE ONLY.

Functional fp32 restatement of ``Conv1DTemporalAttention.forward`` (talking_face/ldm/modules/encoders/modules.py:76-113)
over a plain state dict.  Pinned by ``tests/test_oracle_golden.py`` against the output of the REAL reference module
(``tests/golden/make_golden.py: cond_goldens`` → ``tests/golden/cond.npz``).  The product never imports this module.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# cond_stage_config_2 of mead-128-ldm-f4.yaml:75-80
TF_AUDIO = dict(seq_len=17, subspace_dim=768, subspace2hidden=False)
CHANNELS = [192, 64, 16, 4, 1]      # modules.py:86-96


def cond_param_shapes(cfg: dict) -> Dict[str, Tuple[int, ...]]:
    sh: Dict[str, Tuple[int, ...]] = {}
    cin = cfg["subspace_dim"]
    for i, cout in enumerate(CHANNELS):
        sh[f"attentionConvNet.{2 * i}.weight"] = (cout, cin, 3)
        sh[f"attentionConvNet.{2 * i}.bias"] = (cout,)
        cin = cout
    L = cfg["seq_len"]
    sh["attentionNet.0.weight"] = (L, L)
    sh["attentionNet.0.bias"] = (L,)
    if cfg.get("subspace2hidden"):
        sh["subspace2hidden.weight"] = (cfg["hidden_dim"], cfg["subspace_dim"])
        sh["subspace2hidden.bias"] = (cfg["hidden_dim"],)
    return sh


def init_cond_state_dict(cfg: dict, seed: int = 0) -> Dict[str, Tensor]:
    """Deterministic synthetic weights: gain·sqrt(3)·U(-1,1)/sqrt(fan_in), biases 0.05·N(0,1).  Gains 1.6 (convs: LeakyReLU
    halves the variance per layer) and 4 (Linear) make the attention weights peaked but not one-hot (max 0.26-0.74), so the
    output depends on every conv layer — with unit gains the scores vanish and the softmax is flat."""
    g = torch.Generator().manual_seed(seed)
    sd: Dict[str, Tensor] = {}
    for name, shape in sorted(cond_param_shapes(cfg).items()):
        if len(shape) >= 2:
            fan_in = 1
            for s in shape[1:]:
                fan_in *= s
            gain = 4.0 if name.startswith("attentionNet") else 1.6
            sd[name] = ((torch.rand(shape, generator=g) * 2 - 1) * gain * math.sqrt(3.0 / fan_in)).contiguous()
        else:
            sd[name] = 0.05 * torch.randn(shape, generator=g)
    return sd


@torch.no_grad()
def conv1d_temporal_attention(sd: Dict[str, Tensor], cfg: dict, x: Tensor) -> Tensor:
    """modules.py:101-113.  x [b, seq_len, subspace_dim] -> [b, 1, subspace_dim] (or [b, hidden_dim])."""
    b, L = x.shape[0], cfg["seq_len"]
    xt = torch.transpose(x, 1, 2)
    h = xt
    for i in range(len(CHANNELS)):
        h = F.leaky_relu(F.conv1d(h, sd[f"attentionConvNet.{2 * i}.weight"], sd[f"attentionConvNet.{2 * i}.bias"],
                                  stride=1, padding=1), 0.02)
    att = F.softmax(F.linear(h.view(b, L), sd["attentionNet.0.weight"], sd["attentionNet.0.bias"]), dim=1).view(b, L, 1)
    res = torch.bmm(xt, att).view(b, cfg["subspace_dim"])
    if cfg.get("subspace2hidden"):
        return F.linear(res, sd["subspace2hidden.weight"], sd["subspace2hidden.bias"])
    return res.unsqueeze(1)
