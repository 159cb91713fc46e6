"""Tensor-parallel N-sharding of a quantised Linear (BASELINE.json configs[4]).

Output channels are independent (every row of W with its scale[n], bias[n] produces one output
column, binary/modules.py:104-108), so the layer shards along N with no cross-rank reduction:
each rank owns a contiguous block of rows, runs the single-GPU kernels on its shard with the
replicated activations (the per-token int8 quantisation is identical on every rank, so integer
accumulators do not depend on the world size) and the output shards are exchanged with ONE
collective, an all-gather over NCCL (NVLink 5 / NVSwitch).  The reference has no distributed code.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
from torch import nn

from . import ops

ROW_ALIGN = 16      # the decode GEMV works on 16-row tiles, the tcgen05 GEMM on 128-row tiles


def shard_rows(N: int, world: int, rank: int, align: int = ROW_ALIGN) -> tuple[int, int]:
    """Contiguous row block [start, stop) of rank `rank`: equal blocks of ceil(N/world) rounded up to
    `align`; trailing ranks may get a shorter (or empty) block."""
    per = -(-N // world)
    per = -(-per // align) * align
    start = min(N, rank * per)
    return start, min(N, start + per)


def shard_size(N: int, world: int, align: int = ROW_ALIGN) -> int:
    per = -(-N // world)
    return -(-per // align) * align


class TPShardedLinear(nn.Module):
    """One rank's shard of a Scaled{Ternary,Binary,Binary01}Linear: packed rows, scale and bias of the
    rank's row block.  ``forward(x)`` returns the FULL ``[..., N]`` output on every rank."""

    def __init__(self, packed_shard: torch.Tensor, scale_shard, bias_shard, in_features: int, out_features: int,
                 maxval: float, group=None, matmul=None):
        super().__init__()
        self.in_features, self.out_features, self.maxval = in_features, out_features, maxval
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_shard = shard_size(out_features, self.world)
        self.start, self.stop = shard_rows(out_features, self.world, self.rank)
        self.register_buffer("packed", packed_shard)
        self.register_buffer("scale", scale_shard)
        self.register_buffer("bias", bias_shard)
        self._matmul = matmul or ops.ternary_matmul          # tests inject a CPU stand-in for the gloo path

    @classmethod
    def from_layer(cls, layer, group=None, matmul=None):
        """Shard a full quantised layer (every rank holds the same latent weights, e.g. right after loading
        a checkpoint) by packing only this rank's rows."""
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        a, b = shard_rows(layer.out_features, world, rank)
        w = layer.weights.detach()[a:b]
        packed = ops.quantize_pack(w.contiguous(), layer._mode) if matmul is None else w.contiguous()
        scale = layer.scale.detach()[a:b].float().contiguous() if getattr(layer, "scale", None) is not None else None
        bias = layer.bias.detach()[a:b].float().contiguous() if layer.bias is not None else None
        return cls(packed, scale, bias, layer.in_features, layer.out_features, layer.maxval, group, matmul)

    def local_forward(self, x2: torch.Tensor) -> torch.Tensor:
        """[M, K] -> this rank's output columns, padded to the common shard width [M, n_shard]."""
        M = x2.shape[0]
        out = torch.zeros((M, self.n_shard), dtype=x2.dtype, device=x2.device)
        n_local = self.stop - self.start
        if n_local > 0:
            out[:, :n_local] = self._matmul(x2, self.packed, self.scale, self.bias, self.maxval)
        return out

    def gather(self, y_local: torch.Tensor) -> torch.Tensor:
        """all-gather of the [M, n_shard] shards -> [M, N]."""
        M = y_local.shape[0]
        if self.world == 1:
            return y_local[:, : self.out_features]
        buf = torch.empty((self.world * M, self.n_shard), dtype=y_local.dtype, device=y_local.device)
        dist.all_gather_into_tensor(buf, y_local.contiguous(), group=self.group)      # rank-major concatenation
        return buf.view(self.world, M, self.n_shard).permute(1, 0, 2).reshape(M, self.world * self.n_shard)[:, : self.out_features]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x2 = x.reshape(-1, x.shape[-1])
        y = self.gather(self.local_forward(x2))
        return y.reshape(*x.shape[:-1], self.out_features)
