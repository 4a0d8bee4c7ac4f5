"""Device-time the fused attention kernels at the BERT-base shape against the library SDPA (flash) kernels."""
import json
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200.ops import functional as F  # noqa: E402


def timeit(fn, iters=20):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    B, S, H, d = 32, 512, 12, 64
    q, k, v = ((torch.randn(B, S, H * d, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_() for _ in range(3))
    go = torch.randn(B, S, H * d, device="cuda").to(torch.bfloat16)
    res = {"shape": [B, S, H, d]}
    with torch.no_grad():
        res["fwd_us"] = timeit(lambda: F.attention_fused_fwd(q, k, v, H, want_lse=True))
    out = F.attention(q, k, v, H)
    res["bwd_us"] = timeit(lambda: out.backward(go, retain_graph=True))
    F.dropout_begin_step(q.device)
    with torch.no_grad():
        res["fwd_dropout_us"] = timeit(lambda: F.attention_fused_fwd(q, k, v, H, want_lse=True, drop=(0.1, 1)))
    outd = F.attention(q, k, v, H, None, 0.1, True)
    res["bwd_dropout_us"] = timeit(lambda: outd.backward(go, retain_graph=True))
    sp = lambda t: t.view(B, S, H, d).transpose(1, 2)
    with torch.no_grad():
        res["sdpa_fwd_us"] = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(sp(q), sp(k), sp(v)))
    o2 = torch.nn.functional.scaled_dot_product_attention(sp(q), sp(k), sp(v))
    g2 = sp(go)
    res["sdpa_bwd_us"] = timeit(lambda: o2.backward(g2, retain_graph=True))
    flops = 4 * B * H * S * S * d
    res["fwd_tflops"] = flops / res["fwd_us"] / 1e6
    res["bwd_tflops"] = 2.5 * flops / res["bwd_us"] / 1e6
    print(json.dumps(res))


if __name__ == "__main__":
    main()
