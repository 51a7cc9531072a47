#!/usr/bin/env python
"""tcgen05 GEMM (K10) vs cuBLAS (torch.matmul) on the Linear shapes of GPT-2 medium and the MLP.
CUDA-event timing after warm-up, L2 flushed between iterations; reports TFLOP/s and the fraction of
the measured cuBLAS bf16 peak in MEASURED_PEAKS.json."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from metisfl_b200.ops import gemm  # noqa: E402


def timeit(fn, iters=20, warmup=5):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def gpt2_shapes(tokens=4096, d=1024, vocab=50304):
    """Every GEMM of one GPT-2-medium training step (per layer x 24, head x 1): forward y = x W^T, data gradient
    dx = dy W (B read MN-major), weight gradient dW = dy^T x (both operands MN-major)."""
    out = []
    for name, n_out, n_in, reps in (("qkv", 3 * d, d, 24), ("proj", d, d, 24), ("fc", 4 * d, d, 24), ("out", d, 4 * d, 24),
                                    ("head", vocab, d, 1)):
        out.append((name + ".fwd", tokens, n_out, n_in, False, False, reps))
        out.append((name + ".dgrad", tokens, n_in, n_out, False, True, reps))
        out.append((name + ".wgrad", n_out, n_in, tokens, True, True, reps))
    return out


def gpt2_main():
    rows, tot = [], {"ours": 0.0, "cublas": 0.0, "single": 0.0}
    for name, m, n, k, a_mn, b_mn, reps in gpt2_shapes():
        a = torch.randn((k, m) if a_mn else (m, k), device="cuda").to(torch.bfloat16)
        b = torch.randn((k, n) if b_mn else (n, k), device="cuda").to(torch.bfloat16)
        ref_a = a.t() if a_mn else a
        ref_b = b if b_mn else b.t()
        gemm.set_pair_mode(-1)
        t_ours = timeit(lambda: gemm.gemm_ex(a, b, a_mn, b_mn), iters=10, warmup=3)
        gemm.set_pair_mode(0)
        t_single = timeit(lambda: gemm.gemm_ex(a, b, a_mn, b_mn), iters=10, warmup=3)
        gemm.set_pair_mode(-1)
        t_cublas = timeit(lambda: torch.matmul(ref_a, ref_b), iters=10, warmup=3)
        fl = 2.0 * m * n * k
        rows.append({"gemm": name, "m": m, "n": n, "k": k, "a_mn": a_mn, "b_mn": b_mn, "per_step": reps,
                     "ours_us": t_ours * 1e3, "single_cta_us": t_single * 1e3, "cublas_us": t_cublas * 1e3,
                     "ours_tflops": fl / t_ours / 1e9, "cublas_tflops": fl / t_cublas / 1e9,
                     "ours_over_cublas": t_cublas / t_ours})
        tot["ours"] += t_ours * reps
        tot["cublas"] += t_cublas * reps
        tot["single"] += t_single * reps
    print(json.dumps({"bench": "gemm_bench --gpt2", "step_gemm_ms": tot, "results": rows}))


def main():
    if "--gpt2" in sys.argv:
        return gpt2_main()
    peak = 1642.2
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                           "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        pass
    shapes = [(8192, 3072, 1024), (8192, 1024, 1024), (8192, 4096, 1024), (8192, 1024, 4096), (8192, 8192, 8192),
              (4096, 128, 784), (4096, 128, 128)]
    rows = []
    for m, n, k in shapes:
        a = torch.randn(m, k, device="cuda").to(torch.bfloat16)
        b = torch.randn(n, k, device="cuda").to(torch.bfloat16)
        out = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        gemm.set_pair_mode(-1)
        t_ours = timeit(lambda: gemm.gemm_nt(a, b, out=out))        # default dispatch
        gemm.set_pair_mode(0)
        t_single = timeit(lambda: gemm.gemm_nt(a, b, out=out))      # single-CTA kernel (cta_group::1)
        gemm.set_pair_mode(1)
        t_pair = timeit(lambda: gemm.gemm_nt(a, b, out=out))        # CTA-pair kernel where legal (cta_group::2)
        gemm.set_pair_mode(-1)
        t_cublas = timeit(lambda: torch.matmul(a, b.t(), out=out))
        fl = 2.0 * m * n * k
        rows.append({"m": m, "n": n, "k": k, "ours_ms": t_ours, "cublas_ms": t_cublas,
                     "single_cta_ms": t_single, "cta_pair_ms": t_pair,
                     "ours_tflops": fl / t_ours / 1e9, "cublas_tflops": fl / t_cublas / 1e9,
                     "single_cta_tflops": fl / t_single / 1e9, "cta_pair_tflops": fl / t_pair / 1e9,
                     "ours_over_cublas": t_cublas / t_ours,
                     "ours_fraction_of_measured_peak": fl / t_ours / 1e9 / peak})
    print(json.dumps({"bench": "gemm_bench", "measured_bf16_peak_tflops": peak, "results": rows}))


if __name__ == "__main__":
    main()
