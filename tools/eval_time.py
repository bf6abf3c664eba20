#!/usr/bin/env python
"""Time of ONE UNet eval through the shipped path (CUDA-graph replay of the launch plan), CUDA events around `reps`
replays — the number that matters at small batch, where tools/op_table.py's per-launch events are dominated by the
host's launch cost.  Usage: python tools/eval_time.py --cases 1x64,4x64,1x128 [--reps 30]   (env: LDMB200_SPLITK=0|1 ...)"""
import argparse
import json
import os
import sys

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import dsml_thesis_b200 as P  # noqa: E402
from bench import FR_UNET, init_weights  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="1x64,2x64,4x64,8x64,1x128")
    ap.add_argument("--reps", type=int, default=30)
    ap.add_argument("--precision", default="f16")
    a = ap.parse_args()
    torch.cuda.set_device(0)
    unet = P.UNetModel(**FR_UNET)
    init_weights(unet, 0)
    unet = unet.cuda().eval()
    unet.precision = a.precision
    out = {"split_k": unet.split_k, "precision": a.precision, "cases": {}}
    for case in a.cases.split(","):
        b, hw = (int(v) for v in case.split("x"))
        x = torch.randn(b, 3, hw, hw, device="cuda")
        ctx = torch.randn(b, 1, 512, device="cuda")
        t = torch.full((b,), 501, device="cuda", dtype=torch.long)
        with torch.no_grad():
            for _ in range(4):           # eager, capture, two replays
                unet(x, t, context=ctx)
            torch.cuda.synchronize()
            best = 1e9
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(a.reps):
                    unet(x, t, context=ctx)
                e1.record()
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1) / a.reps)
        out["cases"][case] = round(best, 4)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
