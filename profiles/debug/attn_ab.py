"""Encoder attention of the training path at the c5 size (B = 64, 8 heads, 101 nodes, d = 16, fp32): PyTorch's fused
memory-efficient kernel against the explicit bmm / softmax / bmm statement, forward + backward, both replayed from CUDA
graphs (device time).  usage (GPU box): python profiles/debug/attn_ab.py"""
import math, torch
import torch.nn.functional as F

dev = "cuda"


class Fused(torch.nn.Module):
    def forward(self, q, k, v):
        return F.scaled_dot_product_attention(q, k, v)


class Explicit(torch.nn.Module):
    def forward(self, q, k, v):
        p = torch.softmax(torch.matmul(q, k.transpose(-1, -2)) * (1.0 / math.sqrt(q.shape[-1])), dim=-1)
        return torch.matmul(p, v)


def run(mod, name):
    torch.manual_seed(0)
    q, k, v = (torch.randn(64, 8, 101, 16, device=dev, requires_grad=True) for _ in range(3))
    g = torch.cuda.make_graphed_callables(mod, (q, k, v))
    cot = torch.randn(64, 8, 101, 16, device=dev)
    for _ in range(5):
        g(q, k, v).backward(cot)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        g(q, k, v).backward(cot)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 200 * 1e3:.1f} us per forward + backward")
    return g(q, k, v).detach()


a = run(Fused(), "memory-efficient sdpa")
b = run(Explicit(), "explicit bmm/softmax/bmm")
print("max |diff|", (a - b).abs().max().item())
