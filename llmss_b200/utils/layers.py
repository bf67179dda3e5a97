"""Tensor-parallel layer library (public API parity with the reference).

Parity: reference `src/llmss/server/models/utils/layers.py:12-256` -- `FastLinear`,
`get_linear`, `SuperLayer`, `TensorParallelHead`, `TensorParallelColumnLinear`,
`TensorParallelRowLinear`, `TensorParallelEmbedding`, `LayerNorm.load / load_no_bias`,
`FastLayerNorm`.  Constructors, `load(config, prefix, weights, bias)` classmethods and the
sharding rules are the same, so model code written against the
reference's layers keeps working.  What runs underneath is different:

  * every matmul is the tcgen05 GEMM (ops.linear) instead of cuBLAS `F.linear`;
  * `TensorParallelRowLinear` fuses its all-reduce into the GEMM epilogue over NVLink peer
    memory when a symmetric workspace is attached (otherwise one-shot / NCCL all-reduce);
  * `FastLayerNorm` is always available (own fused residual+LayerNorm kernel) -- the reference
    needs flash-attn's optional `dropout_layer_norm` extension and never instantiates it.

The serving engine does not go through these modules (it uses the fused per-rank layouts in
models/decoder.py); they are the building blocks for user-defined models.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn.functional as F
from torch import nn

from .. import ops
from ..parallel.comm import comm_for


# -- LayerNorm loaders (the reference monkey-patches these onto torch.nn.LayerNorm, layers.py:12-36)
@classmethod
def load_layer_norm(cls, prefix, weights, eps):
    weight = weights.get_tensor(f"{prefix}.weight")
    bias = weights.get_tensor(f"{prefix}.bias")
    with torch.device("meta"):
        ln = cls(weight.shape, eps=eps)
    ln.weight = nn.Parameter(weight, requires_grad=False)
    ln.bias = nn.Parameter(bias, requires_grad=False)
    return ln


@classmethod
def load_layer_norm_no_bias(cls, prefix, weights, eps):
    weight = weights.get_tensor(f"{prefix}.weight")
    with torch.device("meta"):
        ln = cls(weight.shape, eps=eps)
    ln.weight = nn.Parameter(weight, requires_grad=False)
    ln.bias = None
    return ln


torch.nn.LayerNorm.load = load_layer_norm
torch.nn.LayerNorm.load_no_bias = load_layer_norm_no_bias


def _as_2d(x: torch.Tensor):
    return x.reshape(-1, x.shape[-1]), x.shape[:-1]


class FastLinear(nn.Module):
    """y = x @ W^T + b on the tcgen05 GEMM (ref layers.py:39-62)."""

    def __init__(self, weight, bias) -> None:
        super().__init__()
        self.weight = nn.Parameter(weight, requires_grad=False)
        self.bias = nn.Parameter(bias, requires_grad=False) if bias is not None else None

    @classmethod
    def load(cls, config, prefix: str, weights, bias: bool):
        weight = weights.get_tensor(f"{prefix}.weight")
        b = weights.get_tensor(f"{prefix}.bias") if bias else None
        return cls(weight, b)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        x, lead = _as_2d(input)
        if not x.is_cuda and x.dtype == self.weight.dtype:
            return F.linear(input, self.weight, self.bias)
        y = ops.linear(x, self.weight, self.bias)
        return y.reshape(*lead, -1)


def get_linear(weight, bias):
    return FastLinear(weight, bias)


class SuperLayer(nn.Module):
    def __init__(self, linear):
        super().__init__()
        self.linear = linear

    def forward(self, x):
        return self.linear.forward(x)


class TensorParallelHead(SuperLayer):
    """Vocab-parallel LM head + all-gather (ref layers.py:79-135).

    Sharded on dim 0 when ``V % t == 0``; otherwise fully replicated with no gather, exactly
    like the reference (the engine's own head uses ceil-div sharding instead, decoder.py)."""

    def __init__(self, linear, process_group, should_gather: bool):
        super().__init__(linear)
        self.process_group = process_group
        self.should_gather = should_gather

    @staticmethod
    def load(config, prefix: str, weights):
        if weights.process_group.size() > 1:
            try:
                weight = weights.get_sharded(f"{prefix}.weight", dim=0)
                should_gather = True
            except AssertionError:
                weight = weights.get_tensor(f"{prefix}.weight")
                should_gather = False
        else:
            weight = weights.get_tensor(f"{prefix}.weight")
            should_gather = False
        return TensorParallelHead(get_linear(weight, bias=None), process_group=weights.process_group,
                                  should_gather=should_gather)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        out = super().forward(input)
        if not self.should_gather:
            return out
        comm = comm_for(self.process_group, input.device)
        return comm.all_gather_cat(out.contiguous(), dim=-1)


class TensorParallelColumnLinear(SuperLayer):
    """Output-dim sharded linear (ref layers.py:138-153)."""

    @classmethod
    def load(cls, config, prefix: str, weights, bias: bool):
        return cls.load_multi(config, [prefix], weights, bias, dim=0)

    @classmethod
    def load_multi(cls, config, prefixes: List[str], weights, bias: bool, dim: int):
        weight = weights.get_multi_weights_col(prefixes, dim=dim)
        b = torch.cat([weights.get_sharded(f"{p}.bias", dim=0) for p in prefixes], dim=dim) if bias else None
        return cls(get_linear(weight, b))


class TensorParallelRowLinear(SuperLayer):
    """Input-dim sharded linear followed by an all-reduce (ref layers.py:156-179).

    Weight rule and "bias on rank 0 only" rule are the reference's.  On CUDA with a symmetric
    workspace the reduction happens inside the GEMM epilogue (one kernel, K1); the rank-0 bias
    is folded in there after the sum (it is broadcast once at load so every rank adds it once)."""

    def __init__(self, linear, process_group, bias_all=None):
        super().__init__(linear)
        self.process_group = process_group
        # full bias on every rank, used only by the fused epilogue (applied once, after the sum);
        # `linear.bias` keeps the reference rule: present on rank 0 only.
        self.bias_all = bias_all

    @classmethod
    def load(cls, config, prefix: str, weights, bias: bool):
        weight = weights.get_multi_weights_row(prefix)
        b_all = weights.get_tensor(f"{prefix}.bias") if bias else None
        b = b_all if (bias and weights.process_group.rank() == 0) else None  # bias lives on the first rank only
        return cls(get_linear(weight, b), process_group=weights.process_group, bias_all=b_all)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        comm = comm_for(self.process_group, input.device)
        x, lead = _as_2d(input)
        if comm.single:
            return super().forward(input)
        if x.is_cuda and comm.symm is not None:
            y = ops.linear(x.contiguous(), self.linear.weight, self.bias_all, comm=comm)  # fused GEMM -> all-reduce
            return y.reshape(*lead, -1)
        out = super().forward(input)  # rank 0 adds the bias before the sum -> counted once
        comm.all_reduce(out)
        return out


class TensorParallelEmbedding(nn.Module):
    """Vocab-parallel embedding + all-reduce (ref layers.py:182-214): rank block of
    ``V // t`` rows plus one zero row that out-of-shard ids are remapped to."""

    def __init__(self, prefix: str, weights, reduce=True):
        super().__init__()
        weight = weights.get_partial_sharded(f"{prefix}.weight", dim=0)
        num_embeddings = weights.get_shape(f"{prefix}.weight")[0]
        process_group = weights.process_group
        world_size, rank = process_group.size(), process_group.rank()
        block_size = num_embeddings // world_size
        self.min_id = rank * block_size
        self.max_id = min(num_embeddings, (rank + 1) * block_size)
        self.null_idx = block_size
        self.process_group = process_group
        self.reduce = reduce
        self.weight = nn.Parameter(F.pad(weight, (0, 0, 0, 1)), requires_grad=False)

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        idx = torch.where((self.min_id > input) | (input >= self.max_id), self.null_idx, input - self.min_id)
        out = F.embedding(idx, self.weight)
        if self.reduce:
            comm_for(self.process_group, out.device).all_reduce(out)
        return out


class FastLayerNorm(nn.LayerNorm):
    """Fused (residual add +) LayerNorm returning ``(normed, new_residual)`` (ref layers.py:217-256).

    Always available: backed by csrc/norm.cu for hidden <= 8192, PyTorch beyond (same split as
    the reference's `dropout_layer_norm` wrapper)."""

    def forward(self, hidden_states, residual=None):
        if hidden_states.shape[-1] > 8192 or hidden_states.shape[-1] % 8 != 0:
            if residual is not None:
                hidden_states = hidden_states + residual
            return super().forward(hidden_states), hidden_states
        x, lead = _as_2d(hidden_states)
        r = residual.reshape(-1, residual.shape[-1]) if residual is not None else None
        y, new_res = ops.norm_residual(x, r, self.weight, self.bias, self.eps, ops.NORM_LAYER)
        return y.reshape(*lead, -1), new_res.reshape(*lead, -1)
