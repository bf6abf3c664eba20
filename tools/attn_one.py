# one attention_tc launch set at (B, N, heads) for ncu: python tools/attn_one.py [B N h reps]
import sys, torch
sys.path.insert(0, ".")
from dsml_thesis_b200 import _cabi
from dsml_thesis_b200._cabi import F16
B, N, h, reps = (int(a) for a in (sys.argv[1:5] + ["64", "4096", "5", "3"][len(sys.argv) - 1:]))
C = h * 32
torch.manual_seed(0)
qkv = torch.randn(B * N, 3 * C, device="cuda").to(torch.bfloat16)
out = torch.empty(B * N, C, dtype=torch.float16, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(reps):
    _cabi.call("ldmb200_attention_tc", qkv.data_ptr(), 3 * C, 0, C, 2 * C, out.data_ptr(), C, F16, B, h, N, 32 ** -0.5, st)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
