"""ResNet-50 (batch 256) 1x1-convolution GEMM shapes: this library's tcgen05 kernel against torch.matmul (cuBLAS), forward,
data gradient and weight gradient, with the achieved fraction of the memory and tensor rooflines.
usage: python bench/resnet_gemm_shapes.py [out.json]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ml_trainer_b200.ops.gemm import gemm

PEAKS = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))) \
    if os.path.exists("MEASURED_PEAKS.json") else {"hbm_gbs": 6552.0, "bf16_tflops": 1613.7}
SHAPES = [  # (pixels M, Cout N, Cin K)
    (802816, 64, 64), (802816, 256, 64), (802816, 64, 256), (802816, 128, 256), (200704, 512, 128), (200704, 128, 512),
    (200704, 512, 256), (200704, 256, 512), (50176, 1024, 256), (50176, 256, 1024), (50176, 1024, 512), (50176, 512, 1024),
    (12544, 2048, 512), (12544, 512, 2048), (12544, 2048, 1024),
    (802816, 64, 576), (200704, 128, 1152), (50176, 256, 2304), (12544, 512, 4608),
]


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(iters):
                fn()
    best = 1e30
    for _ in range(3):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) * 1e3 / iters)
    return best


rows = []
for M, N, K in SHAPES:
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = torch.randn(N, K, device="cuda").bfloat16()
    dy = torch.randn(M, N, device="cuda").bfloat16()
    wt = w.t().contiguous()
    modes = {
        "fwd": (lambda: gemm(x, w), lambda: torch.matmul(x, w.t()), (M * K + N * K + M * N) * 2),
        "dgrad": (lambda: gemm(dy, w, False, True), lambda: torch.matmul(dy, w), (M * N + N * K + M * K) * 2),
        "wgrad": (lambda: gemm(dy, x, True, True), lambda: torch.matmul(dy.t(), x), (M * N + M * K + N * K) * 2),
    }
    for name, (ours, lib, nbytes) in modes.items():
        t_o, t_l = timed(ours), timed(lib)
        flops = 2.0 * M * N * K
        floor = max(nbytes / (PEAKS["hbm_gbs"] * 1e3), flops / (PEAKS["bf16_tflops"] * 1e6))     # us
        row = dict(M=M, N=N, K=K, mode=name, ours_us=round(t_o, 1), cublas_us=round(t_l, 1), ratio_vs_cublas=round(t_l / t_o, 2),
                   roofline_us=round(floor, 1), frac_of_roofline=round(floor / t_o, 2), gbs=round(nbytes / t_o / 1e3), tflops=round(flops / t_o / 1e6))
        rows.append(row)
        print(json.dumps(row), flush=True)
    del x, w, dy, wt
tot_o = sum(r["ours_us"] for r in rows)
tot_l = sum(r["cublas_us"] for r in rows)
print(json.dumps({"total_ours_us": round(tot_o), "total_cublas_us": round(tot_l), "total_roofline_us": round(sum(r["roofline_us"] for r in rows))}))
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)
