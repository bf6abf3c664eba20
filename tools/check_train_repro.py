#!/usr/bin/env python
"""Run-to-run reproducibility of the training step's gradients: the same step three times from the same state, worst
per-parameter relative-L2 difference between runs.  Usage on the GPU box: python tools/check_train_repro.py [tiny|fr|tf] [hw] [B]"""
import os
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dsml_thesis_b200 as P  # noqa: E402
from dsml_thesis_b200 import train as T  # noqa: E402
from oracle import ldm_oracle as O  # noqa: E402  (configs and seeded weights only)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "tf"
    hw = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    B = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    cfg = {"tiny": O.TINY_UNET, "fr": O.FR_UNET, "tf": O.TF_UNET}[name]
    two = name == "tf"
    unet = P.UNetModel(**cfg)
    unet.load_state_dict(O.init_state_dict(cfg, 0), strict=True)
    unet = unet.cuda().train()
    handle = P.DiffusionHandle(unet, two_cond=two).cuda().train()
    g = torch.Generator().manual_seed(77)
    x0, noise = torch.randn(B, 3, hw, hw, generator=g).cuda(), torch.randn(B, 3, hw, hw, generator=g).cuda()
    ctx = torch.randn(B, 1, cfg["context_dim"], generator=g).cuda()
    c34 = torch.randn(B, 6, hw, hw, generator=g).cuda() if two else None
    t = torch.tensor([731, 12, 400, 999, 0, 250, 600, 3][:B]).cuda()
    runs = []
    for _ in range(3):
        unet.zero_grad(set_to_none=True)
        T.p_losses(handle, x0, ctx, t, noise, c34)[0].backward()
        torch.cuda.synchronize()
        runs.append({n: p.grad.detach().clone() for n, p in unet.named_parameters()})
    worst, wname, nbit = 0.0, "", 0
    for n in runs[0]:
        a = runs[0][n].double()
        if float(a.norm()) <= 1e-6 * a.numel() ** 0.5:
            continue
        for r in runs[1:]:
            e = float((r[n].double() - a).norm() / a.norm())
            nbit += int(torch.equal(r[n], runs[0][n]))
            if e > worst:
                worst, wname = e, n
    print(f"{name} {hw}x{hw} B={B}: worst per-parameter run-to-run rel-L2 {worst:.3e} ({wname}); "
          f"{nbit} of {2 * len(runs[0])} comparisons bit-identical")


if __name__ == "__main__":
    main()
