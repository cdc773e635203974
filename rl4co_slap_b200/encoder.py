"""Graph-attention encoder (stays in PyTorch per the north star).

Runs once per batch before the rollout kernels.  Module / parameter names are laid out so
that ``state_dict()`` keys equal the reference's (``encoder.init_embedding.init_embed.weight``,
``encoder.net.layers.<l>.<0..3>...``), i.e. reference checkpoints load unchanged.

Follows: rl4co/models/zoo/am/encoder.py:14-91, rl4co/models/nn/graph/attnnet.py:16-103,
rl4co/models/nn/attention.py:65-135, rl4co/models/nn/ops.py:12-56, rl4co/models/nn/mlp.py,
rl4co/models/nn/env_embeddings/init.py:51-64,111-134, examples/slap.py:11-47.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


# ---- init embeddings ------------------------------------------------------------------
class TSPInitEmbedding(nn.Module):
    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        return self.init_embed(td["locs"])


class VRPInitEmbedding(nn.Module):
    def __init__(self, embed_dim: int, linear_bias: bool = True, node_dim: int = 3):
        super().__init__()
        self.init_embed = nn.Linear(node_dim, embed_dim, linear_bias)
        self.init_embed_depot = nn.Linear(2, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"]
        depot, cities = locs[:, :1, :], locs[:, 1:, :]
        feats = torch.cat((cities, td["demand"][..., None]), -1)
        return torch.cat((self.init_embed_depot(depot), self.init_embed(feats)), -2)


class SLAPInitEmbedding(nn.Module):
    """x, y and Manhattan distance to node 0 (examples/slap.py:11-47)."""

    def __init__(self, embed_dim: int, linear_bias: bool = True):
        super().__init__()
        self.init_embed = nn.Linear(3, embed_dim, linear_bias)

    def forward(self, td):
        locs = td["locs"].to(torch.float32)
        d0 = (locs - locs[:, :1, :]).abs().sum(-1)
        return self.init_embed(torch.cat((locs, d0[..., None]), -1))


INIT_EMBEDDINGS = {"tsp": TSPInitEmbedding, "cvrp": VRPInitEmbedding, "slap": SLAPInitEmbedding}


# ---- encoder blocks -------------------------------------------------------------------
class _Skip(nn.Module):
    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, x):
        return x + self.module(x)


class _SelfAttention(nn.Module):
    def __init__(self, embed_dim: int, num_heads: int):
        super().__init__()
        self.num_heads = num_heads
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim)
        self.out_proj = nn.Linear(embed_dim, embed_dim)

    def forward(self, x):
        B, N, E = x.shape
        qkv = self.Wqkv(x).view(B, N, 3, self.num_heads, E // self.num_heads)
        q, k, v = qkv.permute(2, 0, 3, 1, 4).unbind(0)
        o = F.scaled_dot_product_attention(q, k, v)
        return self.out_proj(o.transpose(1, 2).reshape(B, N, E))


class _FeedForward(nn.Module):
    def __init__(self, embed_dim: int, hidden: int):
        super().__init__()
        dims = [embed_dim] + ([hidden] if hidden > 0 else []) + [embed_dim]
        self.lins = nn.ModuleList(nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:]))

    def forward(self, x):
        for lin in self.lins[:-1]:
            x = F.relu(lin(x))
        return self.lins[-1](x)


class _Norm(nn.Module):
    def __init__(self, embed_dim: int, kind: str):
        super().__init__()
        self.kind = kind
        if kind == "batch":
            self.normalizer = nn.BatchNorm1d(embed_dim, affine=True)
        elif kind == "instance":
            self.normalizer = nn.InstanceNorm1d(embed_dim, affine=True)
        elif kind == "layer":
            self.normalizer = None
        else:
            raise ValueError(f"unknown normalization {kind!r}")

    def forward(self, x):
        if self.kind == "batch":
            return self.normalizer(x.reshape(-1, x.size(-1))).view(*x.size())
        if self.kind == "instance":
            return self.normalizer(x.permute(0, 2, 1)).permute(0, 2, 1)
        mu = x.mean((1, 2)).view(-1, 1, 1)
        return (x - mu) / torch.sqrt(x.var((1, 2)).view(-1, 1, 1) + 1e-05)


class _Layer(nn.Sequential):
    def __init__(self, embed_dim, num_heads, hidden, normalization):
        super().__init__(
            _Skip(_SelfAttention(embed_dim, num_heads)),
            _Norm(embed_dim, normalization),
            _Skip(_FeedForward(embed_dim, hidden)),
            _Norm(embed_dim, normalization),
        )


class GraphAttentionNetwork(nn.Module):
    def __init__(self, num_heads, embed_dim, num_layers, normalization="batch", feedforward_hidden=512):
        super().__init__()
        self.layers = nn.Sequential(
            *(_Layer(embed_dim, num_heads, feedforward_hidden, normalization) for _ in range(num_layers))
        )

    def forward(self, x, mask=None):
        assert mask is None, "Mask not yet supported!"
        return self.layers(x)


class AttentionModelEncoder(nn.Module):
    def __init__(self, embed_dim: int = 128, init_embedding: nn.Module = None, env_name: str = "tsp",
                 num_heads: int = 8, num_layers: int = 3, normalization: str = "batch",
                 feedforward_hidden: int = 512, net: nn.Module = None, **_):
        super().__init__()
        self.env_name = env_name
        if init_embedding is None:
            if env_name not in INIT_EMBEDDINGS:
                raise ValueError(
                    f"Unknown environment name '{env_name}'. Available init embeddings: {INIT_EMBEDDINGS.keys()}")
            init_embedding = INIT_EMBEDDINGS[env_name](embed_dim)
        self.init_embedding = init_embedding
        self.net = (GraphAttentionNetwork(num_heads, embed_dim, num_layers, normalization, feedforward_hidden)
                    if net is None else net)

    def forward(self, td, mask=None):
        init_h = self.init_embedding(td)
        return self.net(init_h, mask), init_h
