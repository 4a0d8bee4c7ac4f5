"""Device-time the BatchNorm kernels on every distinct (rows, channels) shape of ResNet-50 at batch 256 (cold L2) and print the
achieved bandwidth against the bytes each pass has to move.  Usage: python bench/bn_shapes.py [out.json]"""
import json
import sys

import torch

sys.path.insert(0, ".")
from ml_trainer_b200.ops import conv as C  # noqa: E402

B = 256
SHAPES = [  # (H*W, channels, count in the network, relu, residual)
    (112 * 112, 64, 1, True, False),
    (56 * 56, 64, 6, True, False), (56 * 56, 256, 3, True, True), (56 * 56, 256, 1, False, False),
    (56 * 56, 128, 1, True, False), (28 * 28, 128, 7, True, False), (28 * 28, 512, 4, True, True), (28 * 28, 512, 1, False, False),
    (28 * 28, 256, 1, True, False), (14 * 14, 256, 11, True, False), (14 * 14, 1024, 6, True, True), (14 * 14, 1024, 1, False, False),
    (14 * 14, 512, 1, True, False), (7 * 7, 512, 5, True, False), (7 * 7, 2048, 3, True, True), (7 * 7, 2048, 1, False, False),
]


def timeit(fn, iters=7):
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(2):
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
    rows, tot = [], {"fwd_us": 0.0, "bwd_us": 0.0, "fwd_floor_us": 0.0, "bwd_floor_us": 0.0}
    for hw, ch, count, relu, res in SHAPES:
        M = B * hw
        bn = torch.nn.BatchNorm2d(ch).cuda().to(torch.bfloat16)
        bn.train()
        x = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(B, hw, 1, ch).requires_grad_()
        r = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(B, hw, 1, ch).requires_grad_() if res else None
        go = torch.randn(M, ch, device="cuda").to(torch.bfloat16).view(B, hw, 1, ch)
        with torch.no_grad():
            fwd = timeit(lambda: C.batch_norm_nhwc(x, bn, relu=relu, residual=r))
        y = C.batch_norm_nhwc(x, bn, relu=relu, residual=r)
        def run_bwd():
            x.grad = None                      # a leaf that already holds a gradient would add an accumulation kernel
            if r is not None:
                r.grad = None
            y.backward(go, retain_graph=True)

        bwd = timeit(run_bwd)
        nb = M * ch * 2
        # forward: stats reads x; apply reads x (+ residual), writes y (+ 1/16 mask).  backward: reduce reads dy, x (+ mask);
        # apply reads dy, x (+ mask), writes dx (+ dres)
        fwd_bytes = nb * (3 + (1 if res else 0)) + (nb // 16 if relu else 0)
        bwd_bytes = nb * (5 + (1 if res else 0)) + (2 * (nb // 16) if relu else 0)
        row = {"M": M, "C": ch, "count": count, "relu": relu, "residual": res, "fwd_us": round(fwd, 1), "bwd_us": round(bwd, 1),
               "fwd_GBs": round(fwd_bytes / fwd / 1e3, 0), "bwd_GBs": round(bwd_bytes / bwd / 1e3, 0)}
        rows.append(row)
        tot["fwd_us"] += fwd * count; tot["bwd_us"] += bwd * count
        tot["fwd_floor_us"] += fwd_bytes / 6552e3 * count; tot["bwd_floor_us"] += bwd_bytes / 6552e3 * count
        print(row, flush=True)
        del x, r, go, y
    tot = {k: round(v, 1) for k, v in tot.items()}
    print("network totals (us):", tot)
    if len(sys.argv) > 1:
        json.dump({"batch": B, "rows": rows, "totals_us": tot, "hbm_GBs": 6552}, open(sys.argv[1], "w"), indent=1)


if __name__ == "__main__":
    main()
