"""Per-shape timing of the tcgen05 GEMM against torch.matmul (cuBLAS) for the BERT-base training GEMMs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ml_trainer_b200.ops.gemm import gemm, gemm_batched

DEV = "cuda:0"
T = 16384


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3   # us


def main():
    rows = []
    for (N, K, name) in [(768, 768, "out"), (2304, 768, "qkv"), (3072, 768, "ffn1"), (768, 3072, "ffn2")]:
        x = torch.randn(T, K, device=DEV).bfloat16()
        w = torch.randn(N, K, device=DEV).bfloat16()
        dy = torch.randn(T, N, device=DEV).bfloat16()
        fl = 2.0 * T * N * K
        bias = torch.randn(N, device=DEV).bfloat16()
        aux = torch.empty(T, N, device=DEV, dtype=torch.bfloat16)
        auxk = torch.randn(T, K, device=DEV).bfloat16()
        for tag, ours, ref in [
            ("fwd  ", lambda: gemm(x, w), lambda: torch.matmul(x, w.t())),
            ("fwd+bias", lambda: gemm(x, w, bias=bias), lambda: torch.addmm(bias, x, w.t())),
            ("fwd+gelu+aux", lambda: gemm(x, w, bias=bias, act="gelu", aux_mode=3, aux=aux), None),
            ("fwd+gelu+z", lambda: gemm(x, w, bias=bias, act="gelu", aux_mode=1, aux=aux), None),
            ("dgrad*aux", lambda: gemm(dy, w, False, True, aux_mode=4, aux=auxk), None),
            ("dgrad*gelu'", lambda: gemm(dy, w, False, True, act="gelu", aux_mode=2, aux=auxk), None),
            ("dgrad", lambda: gemm(dy, w, False, True), lambda: torch.matmul(dy, w)),
            ("wgrad", lambda: gemm(dy, x, True, True), lambda: torch.matmul(dy.t(), x)),
            ("wgrad-nosplit", lambda: gemm(dy, x, True, True, split_k=False), None),
        ]:
            to = timeit(ours)
            tr = timeit(ref) if ref else float("nan")
            rows.append((name, tag, to, fl / to / 1e6, tr, fl / tr / 1e6 if ref else float("nan")))
    B, H, S, d = 32, 12, 512, 64
    q = torch.randn(B, S, H * d, device=DEV).bfloat16()
    P = torch.empty(B, H, S, S, device=DEV, dtype=torch.bfloat16)
    o = torch.empty_like(q)
    HD = H * d
    fl = 2.0 * B * H * S * S * d
    t1 = timeit(lambda: gemm_batched(q, HD, d, S * HD, False, q, HD, d, S * HD, False, P, S, S * S, H * S * S, S, S, d, H, B))
    t2 = timeit(lambda: gemm_batched(P, S, S * S, H * S * S, False, q, HD, d, S * HD, True, o, HD, d, S * HD, S, d, S, H, B))
    t3 = timeit(lambda: gemm_batched(P, S, S * S, H * S * S, True, q, HD, d, S * HD, True, o, HD, d, S * HD, S, d, S, H, B))
    rows.append(("attn", "QK^T ", t1, fl / t1 / 1e6, float("nan"), float("nan")))
    rows.append(("attn", "PV   ", t2, fl / t2 / 1e6, float("nan"), float("nan")))
    rows.append(("attn", "P^TdO", t3, fl / t3 / 1e6, float("nan"), float("nan")))
    rows = [r for r in rows if not (r[1] in ("fwd+gelu+aux", "fwd+gelu+z") and r[0] != "ffn1") and not (r[1] in ("dgrad*aux", "dgrad*gelu'") and r[0] != "ffn2")]
    print("%-8s %-14s %10s %10s %10s %10s" % ("layer", "gemm", "ours us", "TFLOP/s", "cublas us", "TFLOP/s"))
    for r in rows:
        print("%-8s %-14s %10.1f %10.1f %10.1f %10.1f" % r)


if __name__ == "__main__":
    main()
