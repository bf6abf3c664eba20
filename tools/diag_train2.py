import sys, torch, time
sys.path.insert(0, ".")
import dsml_thesis_b200 as P
from dsml_thesis_b200 import train as T
from oracle import ldm_oracle as O
cfg = O.TINY_UNET
sd = O.init_state_dict(cfg, 0)
unet = P.UNetModel(**cfg); unet.load_state_dict(sd); unet = unet.cuda().train()
h = P.DiffusionHandle(unet).cuda().train()
g = torch.Generator().manual_seed(77)
x0 = torch.randn(2, 3, 16, 16, generator=g).cuda(); nz = torch.randn(2, 3, 16, 16, generator=g).cuda()
ctx = torch.randn(2, 1, 64, generator=g).cuda(); t = torch.tensor([731, 12]).cuda()
st = None
def run(zero, pre=None):
    if zero: unet.zero_grad(set_to_none=True)
    loss = T.p_losses(h, x0, ctx, t, nz)[0]
    torch.cuda.synchronize()
    if pre: pre()
    torch.cuda.synchronize()
    loss.backward()
    torch.cuda.synchronize()
    return {n: p.grad.clone() for n, p in unet.named_parameters()}
def worst(a, b, scale=1.0):
    w = (0, "")
    for n in a:
        na = float(a[n].norm())
        if na < 1e-5 * a[n].numel() ** 0.5: continue
        e = float((b[n] - scale * a[n]).norm()) / (scale * na)
        if e > w[0]: w = (e, n)
    return w
a = run(True)
st = T.train_state(unet)
print("fresh again      :", worst(a, run(True)))
print("accumulate (x2)  :", worst(a, run(False), 2.0))
print("accumulate (x3)  :", worst(a, run(False), 3.0))
print("manual zero + acc:", worst(a, run(False, lambda: st.flat.zero_())))
def fill(): st.flat.fill_(1.0)
c = run(False, fill)
print("ones + acc       :", worst(a, {n: v - 1.0 for n, v in c.items()}))
b = run(True)
print("---- per-parameter |b-a|/|a| in model order (only output/middle/out)")
for n in a:
    na = float(a[n].norm())
    if na < 1e-5 * a[n].numel() ** 0.5: continue
    e = float((b[n] - a[n]).norm()) / na
    if e > 1e-5:
        print("%.2e %s" % (e, n))
