"""Probe: which summation order reproduces torch's CUDA row norm over 4 elements (F.normalize) bit for bit."""
import itertools
import torch
g = torch.Generator(device="cuda").manual_seed(1)
q = torch.randn(1_000_000, 4, device="cuda", generator=g) * (0.2 + 3 * torch.rand(1_000_000, 1, device="cuda", generator=g))
n_lib = q.norm(dim=1)
sq = q * q
def fma32(a, b, c):
    return (a.double() * b.double() + c.double()).float()
best = []
for p in itertools.permutations(range(4)):
    a, b, c, d = (sq[:, i] for i in p)
    best.append((float((n_lib == torch.sqrt(((a + b) + c) + d)).float().mean()), "seq", p))
    best.append((float((n_lib == torch.sqrt((a + b) + (c + d))).float().mean()), "pair", p))
    x = [q[:, i] for i in p]
    best.append((float((n_lib == torch.sqrt(fma32(x[3], x[3], fma32(x[2], x[2], fma32(x[1], x[1], sq[:, p[0]]))))).float().mean()), "fma", p))
    # two fma chains combined (2 accumulators)
    best.append((float((n_lib == torch.sqrt(fma32(x[1], x[1], sq[:, p[0]]) + fma32(x[3], x[3], sq[:, p[2]]))).float().mean()), "2fma", p))
best.sort(reverse=True)
for b in best[:6]:
    print(b)
# scaled variants (overflow-safe norms): max-scaled
m = q.abs().max(dim=1).values
r = q / m.unsqueeze(1)
print("scaled", float((n_lib == m * torch.sqrt((r * r).sum(1))).float().mean()))
print("sum() of squares", float((n_lib == torch.sqrt(sq.sum(1))).float().mean()))
print("fp64 sum", float((n_lib == q.double().pow(2).sum(1).sqrt().float()).float().mean()))
print("fp32 sqrt of fp64 sum", float((n_lib == torch.sqrt(q.double().pow(2).sum(1).float())).float().mean()))
