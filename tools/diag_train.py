import sys, torch
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
def run(zero):
    if zero: unet.zero_grad(set_to_none=True)
    T.p_losses(h, x0, ctx, t, nz)[0].backward()
    torch.cuda.synchronize()
    return {n: p.grad.clone() for n, p in unet.named_parameters()}
a = run(True); b = run(True); c = run(False)
rows = []
for n in a:
    na = float(a[n].norm())
    if na < 1e-5 * a[n].numel() ** 0.5: continue
    rows.append((float((a[n]-b[n]).norm())/na, float((c[n]-2*a[n]).norm())/(2*na), n))
rows.sort(reverse=True)
for r in rows[:12]: print("%.3e %.3e %s" % r)
print("count a-b > 1e-3:", sum(1 for r in rows if r[0] > 1e-3), "of", len(rows))
print([r[2] for r in rows if r[0] > 1e-3][:60])
_, gr = O.p_losses(sd, cfg, x0.cpu(), t.cpu(), nz.cpu(), ctx.cpu())
for name in ["output_blocks.4.0.emb_layers.1.bias", "output_blocks.5.0.in_layers.2.bias", "output_blocks.5.0.out_layers.3.bias", "out.2.bias", "time_embed.2.bias"]:
    print(name, "a vs ref %.3e  b vs ref %.3e" % (O.rel_l2(a[name].cpu(), gr[name]), O.rel_l2(b[name].cpu(), gr[name])), a[name][:4].tolist(), b[name][:4].tolist(), gr[name][:4].tolist())
print("---- c vs 2a > 5e-4")
for r in sorted(rows, key=lambda r: -r[1]):
    if r[1] > 5e-4: print("%.3e %.3e %s" % r)
