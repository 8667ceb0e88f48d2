"""Probe: per-call time of the fused Adam step back to back, after an unrelated large kernel, and with freshly
allocated gradients (what the training loop does)."""
import torch
from gaussmart_b200.optim import FusedAdam

P = 5_000_000
shapes = [(P, 3), (P, 16, 3), (P, 1), (P, 2), (P, 4)]
params = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in shapes]
for p in params:
    p.grad = torch.randn_like(p) * 1e-3
opt = FusedAdam(params, lr=1e-8, eps=1e-15)
junk = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")


def timed(pre):
    out = []
    for _ in range(6):
        pre()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); opt.step(); e1.record(); torch.cuda.synchronize()
        out.append(round(e0.elapsed_time(e1), 3))
    return out


print("back to back      ", timed(lambda: None))
print("after 1 GB memset ", timed(lambda: junk.zero_()))
print("after sync + sleep", timed(lambda: (torch.cuda.synchronize(), torch.cuda._sleep(2_000_000))))


def fresh():
    for p in params:
        p.grad = None
    for p in params:
        p.grad = torch.randn_like(p) * 1e-3


print("fresh grads       ", timed(fresh))
ref = torch.optim.Adam([torch.nn.Parameter(p.detach().clone()) for p in params], lr=1e-8, eps=1e-15)
for g in ref.param_groups:
    for p in g["params"]:
        p.grad = torch.randn_like(p) * 1e-3
opt = ref
print("torch back to back", timed(lambda: None))
print("torch after memset", timed(lambda: junk.zero_()))

# ---- value dependence: IEEE sqrt / div slow paths
opt = FusedAdam(params, lr=1e-8, eps=1e-15)
for scale, zero_frac in ((1e-3, 0.0), (1e-3, 0.3), (1e-12, 0.0), (1e-17, 0.0), (1e-19, 0.0), (1e-21, 0.0), (0.0, 0.0)):
    for p in params:
        g = torch.randn_like(p) * scale
        if zero_frac:
            g[torch.rand(p.shape[0], device="cuda") < zero_frac] = 0.0
        p.grad = g
        st = opt.state.get(p)
        if st:
            st["exp_avg"].zero_(); st["exp_avg_sq"].zero_()
    print(f"grad scale {scale:g} zero rows {zero_frac}:", timed(lambda: None)[1:])
