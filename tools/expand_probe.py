#!/usr/bin/env python
"""Micro-benchmark + self-check of gsm_sh_grad_expand (the receiving side of the view-parallel exchange) on ONE GPU:
P surfels, V views, a share of zero rows per view; the TMA bulk-store path against the cooperative-copy path
(GSM_NO_BULK=1), same bits required.  Usage: python tools/expand_probe.py [P] [V]"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussmart_b200.rasterizer import sh_grad_expand

P = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
V = int(sys.argv[2]) if len(sys.argv) > 2 else 8
g = torch.Generator(device="cuda").manual_seed(1)
means = torch.randn(P, 3, device="cuda", generator=g)
campos = torch.randn(V, 3, device="cuda", generator=g) * 5
rgb = torch.randn(V, P, 3, device="cuda", generator=g)
rgb[torch.rand(V, P, device="cuda", generator=g) < 0.45] = 0
out = {}
for mode in ("bulk", "coop"):
    if mode == "coop":
        os.environ["GSM_NO_BULK"] = "1"
    else:
        os.environ.pop("GSM_NO_BULK", None)
    for acc in (False, True):
        d = torch.full((P, 16, 3), 0.25 if acc else float("nan"), device="cuda")
        for _ in range(3):
            if acc:
                d.fill_(0.25)
            sh_grad_expand(means, campos, rgb, d, 3, accumulate=acc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 1 if acc else 10
        if acc:
            d.fill_(0.25)
        e0.record()
        for _ in range(reps):
            sh_grad_expand(means, campos, rgb, d, 3, accumulate=acc)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        nbytes = P * (12 + 12 * V + 192 * (2 if acc else 1))
        print(f"{mode} accumulate={acc}: {ms:.3f} ms, {nbytes / ms / 1e6:.0f} GB/s algorithmic")
        out[(mode, acc)] = d.clone()
for acc in (False, True):
    a, b = out[("bulk", acc)], out[("coop", acc)]
    print("accumulate", acc, "identical bits:", torch.equal(a, b), "max abs diff", (a - b).abs().max().item(), "finite", torch.isfinite(a).all().item())
