"""Epilogue cost breakdown on the FFN shapes: same mainloop, different epilogues (one GPU call = all variants)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedb200.ops import cuda_ops as ops
from fedb200.utils.timers import cuda_time_ms

T = int(sys.argv[1]) if len(sys.argv) > 1 else 9187
if len(sys.argv) > 2:
    from fedb200.ops import _lib
    _lib.lib().fedb200_set_gemm_mode(int(sys.argv[2]))
dev = "cuda"
x = torch.randn(T, 768, device=dev).bfloat16()
w1 = (torch.randn(3072, 768, device=dev) * 0.02).bfloat16()
b1 = torch.randn(3072, device=dev)
z = torch.empty(T, 3072, device=dev, dtype=torch.bfloat16)
y = torch.empty(T, 3072, device=dev, dtype=torch.bfloat16)
h = torch.randn(T, 3072, device=dev).bfloat16()
w2 = (torch.randn(768, 3072, device=dev) * 0.02).bfloat16()
b2 = torch.randn(768, device=dev)
res = torch.randn(T, 768, device=dev).bfloat16()
o2 = torch.empty(T, 768, device=dev, dtype=torch.bfloat16)
dy = torch.randn(T, 768, device=dev).bfloat16()
dz = torch.empty(T, 3072, device=dev, dtype=torch.bfloat16)

def t(name, fn, flops):
    ms = cuda_time_ms(fn, warmup=3, iters=20)
    print(f"{name:44s} {ms*1000:7.1f} us  {flops/ms/1e9:7.1f} TFLOP/s")

F = 2 * T * 3072 * 768
t("ffn-up  discard", lambda: ops.gemm(0, x, w1, z, T, 3072, 768, epi=6, bias=b1), F)
t("ffn-up  bias (1 out)", lambda: ops.gemm(0, x, w1, z, T, 3072, 768, epi=ops.EPI_BIAS, bias=b1), F)
t("ffn-up  bias+gelu eval (1 out)", lambda: ops.call("fedb200_gemm_bf16", 0, ops._p(x), 768, ops._p(w1), 768, T, 3072, 768, ops.EPI_BIAS_GELU, None, ops._p(y), 3072, ops._p(b1), None, 0, 0, 0, 0.0, 1, 0, ops._stream()), F)
t("ffn-up  bias+gelu train (2 out)", lambda: ops.gemm(0, x, w1, z, T, 3072, 768, epi=ops.EPI_BIAS_GELU, bias=b1, out2=y), F)
t("ffn-up  torch.matmul", lambda: torch.matmul(x, w1.T, out=z), F)
t("ffn-down discard", lambda: ops.gemm(0, h, w2, o2, T, 768, 3072, epi=6, bias=b2), F)
t("ffn-down bias", lambda: ops.gemm(0, h, w2, o2, T, 768, 3072, epi=ops.EPI_BIAS, bias=b2), F)
t("ffn-down bias+drop+res", lambda: ops.gemm(0, h, w2, o2, T, 768, 3072, epi=ops.EPI_BIAS_DROP_RES, bias=b2, aux=res, seed=3, dropout_p=0.1), F)
t("ffn-down torch.matmul", lambda: torch.matmul(h, w2.T, out=o2), F)
t("dgrad ffn-down discard", lambda: ops.gemm(1, dy, w2, dz, T, 3072, 768, epi=6), F)
t("dgrad ffn-down add(no aux)", lambda: ops.gemm(1, dy, w2, dz, T, 3072, 768, epi=ops.EPI_ADD), F)
t("dgrad ffn-down * gelu'(z)", lambda: ops.gemm(1, dy, w2, dz, T, 3072, 768, epi=ops.EPI_MUL_GELU_GRAD, aux=z), F)
t("dgrad ffn-down torch.matmul", lambda: torch.matmul(dy, w2, out=dz), F)
Fo = 2 * T * 768 * 768
xo = torch.randn(T, 768, device=dev).bfloat16(); wo = (torch.randn(768, 768, device=dev) * 0.02).bfloat16()
t("attn-out discard", lambda: ops.gemm(0, xo, wo, o2, T, 768, 768, epi=6, bias=b2), Fo)
t("attn-out bias+drop+res", lambda: ops.gemm(0, xo, wo, o2, T, 768, 768, epi=ops.EPI_BIAS_DROP_RES, bias=b2, aux=res, seed=3, dropout_p=0.1), Fo)
t("attn-out torch.matmul", lambda: torch.matmul(xo, wo.T, out=o2), Fo)
# pure write bandwidth reference: fill 2 x T x 3072 bf16
t("memset 2 outputs (torch fill)", lambda: (z.fill_(1.0), y.fill_(2.0)), F)
