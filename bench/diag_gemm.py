"""GEMM bring-up diagnostics: prints per-layout error maps instead of asserting (one GPU call = max info)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fedb200.ops import cuda_ops as ops
from fedb200.utils.timers import cuda_time_ms


def run(layout, M, N, K, splits=1):
    torch.manual_seed(0)
    shp = {0: ((M, K), (N, K)), 1: ((M, K), (K, N)), 2: ((K, M), (K, N))}[layout]
    A = torch.randn(*shp[0], device="cuda").bfloat16(); B = torch.randn(*shp[1], device="cuda").bfloat16()
    ref = {0: lambda: A.float() @ B.float().T, 1: lambda: A.float() @ B.float(), 2: lambda: A.float().T @ B.float()}[layout]()
    try:
        if layout == 2:
            out = torch.zeros(M, N, device="cuda")
            ops.gemm(layout, A, B, out, M, N, K, epi=ops.EPI_ATOMIC_F32, splits=splits)
            got = out
        else:
            out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device="cuda")
            ops.gemm(layout, A, B, out, M, N, K, epi=ops.EPI_ADD)
            got = out.float()
        torch.cuda.synchronize()
    except Exception as e:
        print(f"layout {layout} {M}x{N}x{K}: EXC {e}")
        return
    err = (got - ref).abs()
    rel = err.max().item() / (ref.abs().max().item() + 1e-9)
    print(f"layout {layout} {M}x{N}x{K} splits {splits}: max_abs_err {err.max().item():.4f} rel {rel:.5f} nan {torch.isnan(got).sum().item()}")
    if rel > 2e-2 or torch.isnan(got).any():
        bm, bn = min(64, M), min(64, N)
        e = torch.nan_to_num(err, nan=1e9)
        grid = e[: M // bm * bm, : N // bn * bn].reshape(M // bm, bm, N // bn, bn).amax((1, 3))
        print("  bad 64x64 blocks (rows=m, cols=n):")
        for r in range(min(grid.shape[0], 8)):
            print("   ", "".join("X" if v > 0.05 * ref.abs().max().item() else "." for v in grid[r, :32].tolist()))
        # correlation hints: does got match ref with permuted k halves etc.
        print("  got[0,:8]", got[0, :8].tolist())
        print("  ref[0,:8]", ref[0, :8].tolist())


if __name__ == "__main__":
    from fedb200.ops import _lib
    import sys as _sys
    modes = [int(a) for a in _sys.argv[1:] if a.isdigit()] or [0]
    T_list = [9187]
    for layout in ((0, 1, 2) if "--check" in _sys.argv else ()):
        for (M, N, K) in [(128, 64, 64), (128, 128, 64), (128, 256, 64), (128, 256, 256), (256, 512, 128), (304, 768, 768)]:
            run(layout, M, N, K, splits=1 if layout != 2 else 2)
    # timing on BERT-base shapes
    for mode in modes:
      _lib.lib().fedb200_set_gemm_mode(mode)
      for T in T_list:
        print(f"--- gemm mode {mode} (0 auto/pairs, 1 single-CTA), tokens {T}")
        for name, layout, M, N, K, epi, splits in [("fwd qkv NO-EPILOGUE", 0, T, 2304, 768, 6, 1), ("fwd ffn-down NO-EPILOGUE", 0, T, 768, 3072, 6, 1),
                                                   ("fwd ffn-up NO-EPILOGUE", 0, T, 3072, 768, 6, 1), ("fwd qkv", 0, T, 2304, 768, ops.EPI_BIAS, 1), ("fwd ffn-up", 0, T, 3072, 768, ops.EPI_BIAS, 1),
                                                   ("fwd ffn-down", 0, T, 768, 3072, ops.EPI_BIAS, 1), ("dgrad ffn-down", 1, T, 3072, 768, ops.EPI_ADD, 1),
                                                   ("fwd attn-out", 0, T, 768, 768, ops.EPI_BIAS, 1), ("dgrad qkv", 1, T, 768, 2304, ops.EPI_ADD, 1),
                                                   ("wgrad ffn-up", 2, 3072, 768, T, ops.EPI_ATOMIC_F32, 2), ("wgrad attn-out", 2, 768, 768, T, ops.EPI_ATOMIC_F32, 8)]:
            shp = {0: ((M, K), (N, K)), 1: ((M, K), (K, N)), 2: ((K, M), (K, N))}[layout]
            A = torch.randn(*shp[0], device="cuda").bfloat16(); B = torch.randn(*shp[1], device="cuda").bfloat16()
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if layout == 2 else torch.bfloat16)
            try:
                ms = cuda_time_ms(lambda: ops.gemm(layout, A, B, out, M, N, K, epi=epi, splits=splits), warmup=3, iters=10)
                a2, b2 = (A, B.T) if layout == 0 else ((A, B) if layout == 1 else (A.T, B))
                ms_ref = cuda_time_ms(lambda: torch.matmul(a2, b2), warmup=3, iters=10)
                print(f"time {name}: {ms:.3f} ms = {2*M*N*K/ms/1e9:.1f} TFLOP/s   (torch.matmul {ms_ref:.3f} ms = {2*M*N*K/ms_ref/1e9:.1f} TFLOP/s)")
            except Exception as e:
                print(f"time {name}: EXC {e}")
