import os, sys, subprocess, json
# A/B of attention_tc variants: each library build in its own process (LDMB200_LIB), same inputs
code = r'''
import sys, torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import kernel_harness as K
from dsml_thesis_b200 import _cabi
from dsml_thesis_b200._cabi import BF16, F16
torch.manual_seed(0)
for (B, N, h) in [(64, 4096, 5), (64, 1024, 10), (8, 16384, 5)]:
    C = h * 32
    qkv = (torch.randn(B * N, 3 * C, device="cuda") * 1.0).to(torch.bfloat16)
    out = torch.empty(B * N, C, dtype=torch.float16, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    f = lambda: _cabi.call("ldmb200_attention_tc", qkv.data_ptr(), 3 * C, 0, C, 2 * C, out.data_ptr(), C, F16, B, h, N, 32 ** -0.5, st)
    for _ in range(3): f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    # accuracy on a slice
    q, k, v = (qkv[:N, i * C:(i + 1) * C].float().reshape(1, N, C) for i in range(3))
    ref = K.attention_reference(q, k, v, h, 32 ** -0.5)
    err = K.rel_l2(out[:N].float().reshape(1, N, C), ref)
    print(f"B{B} N{N} h{h}: {ms:.4f} ms  {4.0*B*h*N*N*32/ms/1e9:.1f} TF/s  rel-L2 {err:.2e}")
'''
for k in sys.argv[1:]:
    env = dict(os.environ, LDMB200_LIB=os.path.abspath(f"dsml_thesis_b200/libldmb200_p{k}.so"))
    print("== variant", k); sys.stdout.flush()
    subprocess.run([sys.executable, "-c", code], env=env)
