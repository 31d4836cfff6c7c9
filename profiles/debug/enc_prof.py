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
x = torch.randn(B, n, 128, device=dev); f = torch.randn(B, n, 512, device=dev)
L = _lib.lib()
out = (C.c_longlong * 16)()
names = ["issue/other until done-wait", "wait done (MMA tail)", "epilogue", "end sync"]
with torch.no_grad():
    for name, fn in [("qkv", lambda: net._linear(x, mha.Wqkv, 0, 384)), ("out+norm", lambda: net._linear(x, mha.out_proj, 2, 128, residual=x, norm=norm1)),
                     ("ffn1", lambda: net._linear(x, mlp.lins[0], 1, 512)), ("ffn2+norm", lambda: net._linear(f, mlp.lins[1], 2, 128, residual=x, norm=norm2))]:
        fn(); torch.cuda.synchronize()
        L.rl4co_enc_debug_prof(out, 1)
        fn(); L.rl4co_enc_debug_prof(out, 0)
        inst = (B + 147) // 148
        print(name, "cycles per instance (epilogue thread 0):", {names[i]: out[i] // inst for i in range(4)})
