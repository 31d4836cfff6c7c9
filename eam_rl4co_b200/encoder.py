"""Graph-attention encoder of the AttentionModel, plain PyTorch.

NOT part of the hot path: it runs once per rollout and produces
the node embeddings `h [B, N', E]` the rollout kernels consume.  It is restated here only so
that the policy is usable stand-alone; module / parameter names follow the reference
(rl4co/models/zoo/am/encoder.py:70-91, nn/graph/attnnet.py:16-103, nn/attention.py:66-136,
nn/ops.py:32-56, nn/mlp.py, nn/env_embeddings/init.py:55-68,115-138) so that reference
state_dicts load unchanged.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class TSPInitEmbedding(nn.Module):
    def __init__(self, embed_dim):
        super().__init__()
        self.init_embed = nn.Linear(2, embed_dim, True)

    def forward(self, td):
        return self.init_embed(td["locs"])


class VRPInitEmbedding(nn.Module):
    def __init__(self, embed_dim):
        super().__init__()
        self.init_embed = nn.Linear(3, embed_dim, True)
        self.init_embed_depot = nn.Linear(2, embed_dim, True)

    def forward(self, td):
        depot, cities = td["locs"][:, :1, :], td["locs"][:, 1:, :]
        return torch.cat((self.init_embed_depot(depot),
                          self.init_embed(torch.cat((cities, td["demand"][..., None]), -1))), -2)


class MultiHeadAttention(nn.Module):
    def __init__(self, embed_dim, num_heads):
        super().__init__()
        self.num_heads = num_heads
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim, bias=True)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=True)

    def forward(self, x):
        B, S, E = x.shape
        qkv = self.Wqkv(x).view(B, S, 3, self.num_heads, E // self.num_heads)
        q, k, v = qkv.permute(2, 0, 3, 1, 4).unbind(0)
        out = F.scaled_dot_product_attention(q, k, v)
        return self.out_proj(out.transpose(1, 2).reshape(B, S, E))


class SkipConnection(nn.Module):
    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, x):
        return x + self.module(x)


class Normalization(nn.Module):
    def __init__(self, embed_dim, normalization="batch"):
        super().__init__()
        cls = {"batch": nn.BatchNorm1d, "instance": nn.InstanceNorm1d}[normalization]
        self.normalizer = cls(embed_dim, affine=True)

    def forward(self, x):
        if isinstance(self.normalizer, nn.BatchNorm1d):
            return self.normalizer(x.reshape(-1, x.size(-1))).view(*x.size())
        # nn/ops.py:48-49: InstanceNorm1d over the node axis of [B, N, E].  Same statistics (biased variance,
        # eps inside the sqrt, affine), computed in place on the [B, N, E] layout: the permute -> cuDNN
        # batch-norm route of nn.InstanceNorm1d costs ~1 ms per call at B=1024 on a B200, 12 calls per encode.
        nrm = self.normalizer
        mean = x.mean(dim=1, keepdim=True)
        xc = x - mean
        var = (xc * xc).mean(dim=1, keepdim=True)
        return xc * torch.rsqrt(var + nrm.eps) * nrm.weight + nrm.bias


class MLP(nn.Module):
    def __init__(self, embed_dim, hidden):
        super().__init__()
        self.lins = nn.ModuleList([nn.Linear(embed_dim, hidden), nn.Linear(hidden, embed_dim)])

    def forward(self, x):
        return self.lins[1](F.relu(self.lins[0](x)))


class GraphAttentionNetwork(nn.Module):
    def __init__(self, num_heads, embed_dim, num_layers, normalization, feedforward_hidden):
        super().__init__()
        self.layers = nn.Sequential(*(
            nn.Sequential(
                SkipConnection(MultiHeadAttention(embed_dim, num_heads)),
                Normalization(embed_dim, normalization),
                SkipConnection(MLP(embed_dim, feedforward_hidden)),
                Normalization(embed_dim, normalization),
            ) for _ in range(num_layers)))

    def forward(self, x):
        return self.layers(x)


class AttentionModelEncoder(nn.Module):
    def __init__(self, embed_dim=128, env_name="tsp", num_heads=8, num_layers=3,
                 normalization="batch", feedforward_hidden=512):
        super().__init__()
        self.init_embedding = (TSPInitEmbedding if env_name == "tsp" else VRPInitEmbedding)(embed_dim)
        self.net = GraphAttentionNetwork(num_heads, embed_dim, num_layers, normalization,
                                         feedforward_hidden)

    def forward(self, td):
        init_h = self.init_embedding(td)
        return self.net(init_h), init_h
