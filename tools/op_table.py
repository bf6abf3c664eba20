#!/usr/bin/env python
"""Per-launch table of one UNet eval (CUDA events around every C-ABI launch): kernel, shape, ms, TFLOP/s or GB/s.
Usage on the GPU box: python tools/op_table.py [--batch 64] [--hw 64] [--reps 3] > gpurun_out/op_table.txt"""
import argparse
import os
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dsml_thesis_b200 as P  # noqa: E402
from bench import FR_UNET, init_weights  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--hw", type=int, default=64)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--precision", default="f16")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    unet = P.UNetModel(**FR_UNET)
    init_weights(unet, 0)
    unet = unet.cuda().eval()
    unet.precision = a.precision
    x = torch.randn(a.batch, 3, a.hw, a.hw, device="cuda")
    ctx = torch.randn(a.batch, 1, 512, device="cuda")
    t = torch.full((a.batch,), 501, device="cuda", dtype=torch.long)
    with torch.no_grad():
        unet(x, t, context=ctx)
        plan = next(iter(unet._plans.values()))
        st = torch.cuda.current_stream()
        runs = [plan.run_timed(st) for _ in range(a.reps + 1)][1:]
    n = len(runs[0])
    rows = []
    for i in range(n):
        ms = min(r[i][1] for r in runs)
        name, _, fl, by, what = runs[0][i]
        rows.append((i, name.replace("ldmb200_", ""), what, ms, fl, by))
    tot = sum(r[3] for r in rows)
    print(f"# one UNet eval, B={a.batch}, {a.hw}x{a.hw}, {a.precision}: {n} launches, sum of kernel times {tot:.3f} ms")
    print(f"{'idx':>4} {'kernel':<20} {'shape':<44} {'ms':>8} {'TFLOP/s':>8} {'GB/s':>8}")
    for i, name, what, ms, fl, by in rows:
        tf = fl / (ms * 1e-3) / 1e12 if fl else 0
        gb = by / (ms * 1e-3) / 1e9 if by else 0
        print(f"{i:>4} {name:<20} {what:<44} {ms:8.4f} {tf:8.1f} {gb:8.0f}")
    # aggregate by (kernel, shape)
    agg = {}
    for i, name, what, ms, fl, by in rows:
        k = (name, what)
        v = agg.setdefault(k, [0, 0.0, 0.0, 0.0])
        v[0] += 1; v[1] += ms; v[2] += fl; v[3] += by
    print("\n# aggregated by (kernel, shape), sorted by time")
    print(f"{'kernel':<20} {'shape':<44} {'n':>3} {'ms':>8} {'share':>6} {'TFLOP/s':>8} {'GB/s':>8}")
    for (name, what), (c, ms, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        tf = fl / (ms * 1e-3) / 1e12 if fl else 0
        gb = by / (ms * 1e-3) / 1e9 if by else 0
        print(f"{name:<20} {what:<44} {c:>3} {ms:8.3f} {100*ms/tot:5.1f}% {tf:8.1f} {gb:8.0f}")


if __name__ == "__main__":
    main()
