import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib
import bench
B, n = 1024, 100
dev = torch.device("cuda", 0)
pol = bench.make_policy().to(dev)
net = pol.encoder.net
layer = net.layers[0]
mha, norm1, mlp, norm2 = layer[0].module, layer[1], layer[2].module, layer[3]
x = torch.randn(B, n, 128, device=dev)
L = _lib.lib()
out = (C.c_longlong * 16)()
names = ["issuer wait W tile", "issuer (unused)", "issuer wait a_ready", "issuer wait wfree", "issuer issue", "issuer tail waits",
         "-", "final epilogue + next X (thread 0)", "epi wait acc_full", "conv (unused)", "epi drain/relu/split/store"]
with torch.no_grad():
    net._ffn(x, mlp, norm2); torch.cuda.synchronize()
    L.rl4co_enc_debug_prof(out, 1)
    net._ffn(x, mlp, norm2); L.rl4co_enc_debug_prof(out, 0)
    inst = (B + 147) // 148
    for i, nm in enumerate(names):
        print(f"{nm:40s} {out[i] // inst:8d} cycles per instance")
