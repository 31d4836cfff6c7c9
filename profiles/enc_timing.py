"""CUDA-event timing of the encoder kernels at the bench shape (B=1024, n=100).  usage: python profiles/enc_timing.py"""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib
import bench

def timed(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

def main():
    B, n = 1024, 100
    dev = torch.device("cuda", 0)
    pol = bench.make_policy().to(dev)
    net = pol.encoder.net
    layer = net.layers[0]
    mha, norm1, mlp, norm2 = layer[0].module, layer[1], layer[2].module, layer[3]
    x = torch.randn(B, n, 128, device=dev)
    f = torch.randn(B, n, 512, device=dev)
    qkv = torch.randn(B, n, 384, device=dev)
    heads = torch.empty(B, n, 128, device=dev)
    L = _lib.lib()
    with torch.no_grad():
        print(f"qkv   linear 128->384        {timed(lambda: net._linear(x, mha.Wqkv, 0, 384)):8.3f} ms")
        print(f"attention                    {timed(lambda: L.rl4co_enc_attention(_lib.ptr(qkv), _lib.ptr(heads), C.c_int64(B), n, _lib.stream_ptr())):8.3f} ms")
        print(f"out   linear 128->128 + norm {timed(lambda: net._linear(x, mha.out_proj, 2, 128, residual=x, norm=norm1)):8.3f} ms")
        print(f"ffn1  linear 128->512 relu   {timed(lambda: net._linear(x, mlp.lins[0], 1, 512)):8.3f} ms")
        print(f"ffn2  linear 512->128 + norm {timed(lambda: net._linear(f, mlp.lins[1], 2, 128, residual=x, norm=norm2)):8.3f} ms")
        print(f"fused ffn (ffn1+ffn2+norm)   {timed(lambda: net._ffn(x, mlp, norm2)):8.3f} ms")
        print(f"fused attention block        {timed(lambda: net._attn_block(x, mha, norm1)):8.3f} ms")
        td = P.TSPEnv(generator_params=dict(num_loc=n), device=dev).reset(P.TensorDict({"locs": bench.make_instances(B).to(dev)}, batch_size=[B]))
        print(f"encoder, kernels             {timed(lambda: pol.encoder(td)):8.3f} ms")
        net.fuse_attention = False
        print(f"encoder, kernels, unfused attention {timed(lambda: pol.encoder(td)):8.3f} ms")
        net.fuse_ffn = False
        print(f"encoder, kernels, 5 launches/layer  {timed(lambda: pol.encoder(td)):8.3f} ms")
        net.use_kernels = False
        print(f"encoder, torch               {timed(lambda: pol.encoder(td)):8.3f} ms")

if __name__ == "__main__":
    main()
