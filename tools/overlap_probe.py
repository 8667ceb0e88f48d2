#!/usr/bin/env python
"""Does a kernel on a side stream run NEXT TO a kernel on the current stream?  (single GPU)
A one-CTA spin kernel (torch.cuda._sleep, ~2 ms) on the current stream, the SH expand (39 k CTAs) on a side stream:
the expand's end time tells whether it ran under the spin or after it."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussmart_b200.rasterizer import sh_grad_expand

P, V = 5_000_000, 2
g = torch.Generator(device="cuda").manual_seed(1)
means = torch.randn(P, 3, device="cuda", generator=g)
campos = torch.randn(V, 3, device="cuda", generator=g) * 5
rgb = torch.randn(V, P, 3, device="cuda", generator=g)
d = torch.empty(P, 16, 3, device="cuda")
side = torch.cuda.Stream()
main = torch.cuda.current_stream()
print("main stream:", main, "side:", side, "CUDA_DEVICE_MAX_CONNECTIONS =", os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"))
for order in ("side-first", "main-first"):
    for _ in range(2):
        sh_grad_expand(means, campos, rgb, d, 3)
    torch.cuda.synchronize()
    e0, e_main, e_side = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(main)
    side.wait_stream(main)
    if order == "side-first":
        with torch.cuda.stream(side):
            sh_grad_expand(means, campos, rgb, d, 3)
            e_side.record(side)
        torch.cuda._sleep(int(2e-3 * 1.9e9))
        e_main.record(main)
    else:
        torch.cuda._sleep(int(2e-3 * 1.9e9))
        e_main.record(main)
        with torch.cuda.stream(side):
            sh_grad_expand(means, campos, rgb, d, 3)
            e_side.record(side)
    torch.cuda.synchronize()
    print(order, "spin on main ends at %.3f ms, expand on side ends at %.3f ms" % (e0.elapsed_time(e_main), e0.elapsed_time(e_side)))
