import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn.functional as F
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib
torch.manual_seed(0)
pol = P.POMOPolicy("tsp").eval().cuda()
net = pol.encoder.net
mha = net.layers[0][0].module
B, n = 2, 100
x = torch.randn(B, n, 128, device="cuda")
with torch.no_grad():
    got = net._linear(x, mha.Wqkv, _lib.ENC_EPI_PLAIN, 384)
    ref = F.linear(x.double(), mha.Wqkv.weight.double(), mha.Wqkv.bias.double()).float()
torch.cuda.synchronize()
d = (got - ref).abs()
print("max err", d.max().item(), "ref max", ref.abs().max().item())
print("err by instance", d.amax((1, 2)).tolist())
print("err by 128-col block", [d[:, :, i*128:(i+1)*128].max().item() for i in range(3)])
print("err by row group of 8 (inst 0, block 0):", [round(d[0, i*8:(i+1)*8, :128].max().item(), 3) for i in range(13)])
print("err by 16-col group (inst 0, block 0):", [round(d[0, :, i*16:(i+1)*16].max().item(), 3) for i in range(8)])
print("got[0,0,:8]", got[0,0,:8].tolist()); print("ref[0,0,:8]", ref[0,0,:8].tolist())
# without bias
gb = got - mha.Wqkv.bias
rb = ref - mha.Wqkv.bias
print("ratio got/ref [0,0,:8]", (gb[0,0,:8] / rb[0,0,:8]).tolist())
