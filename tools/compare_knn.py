#!/usr/bin/env python
"""GPU-box tool: distCUDA2 of gaussmart_b200 vs the reference build (bit-exact) + timing."""
import argparse, json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import gaussmart_b200 as gb
from gaussmart_b200 import synthetic
import ref_loader


def timed(fn, iters=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


ap = argparse.ArgumentParser()
ap.add_argument("--P", type=int, nargs="+", default=[10000, 1000000])
ap.add_argument("--out", default=None)
ap.add_argument("--dups", action="store_true")
a = ap.parse_args()
ref = ref_loader.load_knn()
for P in a.P:
    scene = synthetic.make_scene(P, 1237, 0, 4242 + P)
    pts = scene.means3D.cuda()
    if a.dups:
        pts[: P // 10] = pts[P // 2: P // 2 + P // 10]
    r = ref.distCUDA2(pts); n = gb.distCUDA2(pts)
    torch.cuda.synchronize()
    res = dict(P=P, mismatch=int((r.view(torch.int32) != n.view(torch.int32)).sum().item()),
               max_abs=(r - n).abs().max().item(), ms_ref=timed(lambda: ref.distCUDA2(pts)),
               ms_new=timed(lambda: gb.distCUDA2(pts)))
    s = json.dumps(res); print(s)
    if a.out:
        open(a.out, "a").write(s + "\n")
