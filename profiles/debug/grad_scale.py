"""Relative error of the fused evaluate gradients against fp64 as the cotangent shrinks (the split-fp16 operands of the backward
MMAs have fp16's exponent range): usage (GPU box) python profiles/debug/grad_scale.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import torch
from eam_rl4co_b200 import train
from test_train_gpu import _eager_rows

DEV = "cuda"
B, R, n, clip, temp = 2, 300, 101, 10.0, 1.0
g = torch.Generator().manual_seed(5)
rnd = lambda *s: torch.randn(*s, generator=g)
q, gk, gv, lkf = rnd(B, R, 128), rnd(B, n, 128), rnd(B, n, 128), rnd(B, n, 128) * 0.3
mask = torch.rand(B, R, n, generator=g) < 0.5
taken = torch.randint(0, n, (B, R), generator=g)
mask.scatter_(2, taken[..., None], True)
cot0 = rnd(B, R)
for scale in (1.0, 1e-2, 1e-4, 1e-6, 1e-8):
    cot = cot0 * scale
    a64 = [t.double().to(DEV).requires_grad_(True) for t in (q, gk, gv, lkf)]
    ref = _eager_rows(*a64, mask.to(DEV), taken.to(DEV), clip, temp)
    (ref * cot.double().to(DEV)).sum().backward()
    a32 = [t.to(DEV).requires_grad_(True) for t in (q, gk, gv, lkf)]
    got = train._FusedEvaluate.apply(*a32, mask.to(DEV), taken.to(DEV), clip, temp)
    (got * cot.to(DEV)).sum().backward()
    errs = []
    for x32, x64 in zip(a32, a64):
        sc = max(x64.grad.abs().max().item(), 1e-300)
        errs.append((x32.grad.double() - x64.grad).abs().max().item() / sc)
    print(f"cotangent scale {scale:8.0e}: rel err dq {errs[0]:.2e} dgk {errs[1]:.2e} dgv {errs[2]:.2e} dlkf {errs[3]:.2e}")
