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

import ctypes

from . import _lib as L
from . import ops
from ._lib import TernaryB200Error

ROW_ALIGN = 16      # the decode GEMV works on 16-row tiles, the tcgen05 GEMM on 128-row tiles


def shard_rows(N: int, world: int, rank: int, align: int = ROW_ALIGN) -> tuple[int, int]:
    """Contiguous row block [start, stop) of rank `rank`: equal blocks of ceil(N/world) rounded up to
    `align`; trailing ranks may get a shorter (or empty) block."""
    per = -(-N // world)
    per = -(-per // align) * align
    start = min(N, rank * per)
    return start, min(N, start + per)


def _rotated(rank: int, world: int) -> list[int]:
    """Peer order of a rank's stores: rank+1, rank+2, ..., rank (an all-to-all schedule: at any moment the ranks
    address different receivers instead of all starting with rank 0)."""
    return [(rank + 1 + i) % world for i in range(world)]


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
        if matmul is not None:
            packed = w.contiguous()
        elif b > a:
            packed = ops.quantize_pack(w.contiguous(), layer._mode)
        else:       # empty trailing shard (N not a multiple of world x 16): nothing to pack, the rank only gathers
            packed = torch.empty((0, ops.packed_pitch(layer.in_features)), dtype=torch.uint8, device=w.device)
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

    # -- fused gather over NVLink peer memory (decode shapes) -------------------------------------------
    def enable_peer_gather(self, max_tokens: int = 32, dtype=torch.bfloat16, multicast="auto"):
        """Allocate two symmetric (peer-mapped) output buffers [max_tokens, world*n_shard] and exchange their
        addresses.  Afterwards ``forward`` for <= ``max_tokens`` tokens runs ``tb200_linear_forward_tp``: the GEMV
        epilogue (<= 32 tokens) or the tcgen05 GEMM's epilogue (prefill: TMA tile stores) writes this rank's column block
        into every rank's buffer, then ONE cross-rank barrier replaces the NCCL all-gather.  Buffers alternate per call, so a peer may start the next call while this rank's
        consumers still read the previous result."""
        import torch.distributed._symmetric_memory as symm_mem

        dev = self.packed.device
        self._peer_ld = self.world * self.n_shard
        self._peer_tokens = max_tokens
        self._peer_dtype = dtype
        grp = self.group if self.group is not None else dist.group.WORLD
        self._peer_bufs, self._peer_hdls, self._peer_ptrs = [], [], []
        esz = torch.empty((), dtype=dtype).element_size()
        for _ in range(2):
            buf = symm_mem.empty((max_tokens, self._peer_ld), dtype=dtype, device=dev)
            hdl = symm_mem.rendezvous(buf, grp)
            # destinations in rotated order (rank+1, rank+2, ..., rank): the ranks do not all store to the same peer first
            ptrs = (ctypes.c_void_p * self.world)(*[int(hdl.buffer_ptrs[r]) + self.rank * self.n_shard * esz
                                                    for r in _rotated(self.rank, self.world)])
            self._peer_bufs.append(buf)
            self._peer_hdls.append(hdl)
            self._peer_ptrs.append(ptrs)
        self._peer_mc = [_multicast_base(h, self.rank * self.n_shard * esz) for h in self._peer_hdls]
        want_mc = multicast is True or multicast == "try" or (multicast == "auto" and self.world > 4)        # see TPShardedStream
        self._peer_use_mc = bool(want_mc and self.world > 1 and all(self._peer_mc) and (self.stop - self.start) % (16 // esz) == 0)
        self._peer_turn = 0
        return self

    def forward_peer(self, x2: torch.Tensor, alias_ok: bool = False) -> torch.Tensor:
        """Fused peer gather.  Every rank MUST take this path together (the barrier is collective): a rank whose row
        block is empty (N not a multiple of world x 16) skips the kernel launch but still takes the barrier.
        The result lives in one of two internal symmetric buffers that peers overwrite TWO calls later; by default
        a copy is returned (tiny for <= 32 tokens).  ``alias_ok=True`` returns the view itself: valid only until the
        second following call on this layer."""
        M, K = x2.shape
        i = self._peer_turn
        self._peer_turn ^= 1
        buf, hdl, ptrs = self._peer_bufs[i], self._peer_hdls[i], self._peer_ptrs[i]
        n_local = self.stop - self.start
        if n_local > 0 and self._peer_use_mc and M > L.GEMV_MAX_M:
            # prefill: every output tile leaves this GPU once, the NVSwitch replicates it into all ranks' buffers
            ws = ops._workspace(M, K, x2.device)
            with torch.cuda.device(x2.device):
                L.check(L.lib().tb200_linear_forward_mc(
                    x2.data_ptr(), ops._dt(x2), M, K, x2.stride(0), self.packed.data_ptr(), self.packed.stride(0), n_local,
                    ops._ptr(self.scale), ops._ptr(self.bias), float(self.maxval), self._peer_mc[i],
                    ops._DT[self._peer_dtype], self._peer_ld, ws.data_ptr(), ops._stream(x2)), "linear_forward_mc")
        elif n_local > 0:
            ws = ops._workspace(M, K, x2.device)
            with torch.cuda.device(x2.device):
                L.check(L.lib().tb200_linear_forward_tp(
                    x2.data_ptr(), ops._dt(x2), M, K, x2.stride(0), self.packed.data_ptr(), self.packed.stride(0), n_local,
                    ops._ptr(self.scale), ops._ptr(self.bias), float(self.maxval), ptrs, self.world,
                    ops._DT[self._peer_dtype], self._peer_ld, ws.data_ptr(), ops._stream(x2)), "linear_forward_tp")
        hdl.barrier(channel=0)                   # every rank's block has landed in every buffer
        y = buf[:M, : self.out_features]
        return y if alias_ok else y.clone()

    def uses_peer_gather(self, x2: torch.Tensor) -> bool:
        """Path choice of ``forward``: depends only on rank-uniform facts (never on this rank's shard size), so all
        ranks issue the same collective."""
        return bool(getattr(self, "_peer_bufs", None)) and x2.shape[0] <= self._peer_tokens \
            and x2.is_cuda and x2.dtype == self._peer_dtype

    def forward(self, x: torch.Tensor, alias_ok: bool = False) -> torch.Tensor:
        """Full ``[..., N]`` output on every rank.  <= 32 tokens with ``enable_peer_gather``: the fused NVLink gather
        (returns a fresh tensor unless ``alias_ok``, see ``forward_peer``); otherwise NCCL ``all_gather_into_tensor``."""
        x2 = x.reshape(-1, x.shape[-1])
        if self.uses_peer_gather(x2):
            y = self.forward_peer(x2 if x2.stride(-1) == 1 else x2.contiguous(), alias_ok)
        else:
            y = self.gather(self.local_forward(x2))
        return y.reshape(*x.shape[:-1], self.out_features)


def _multicast_base(hdl, offset_bytes: int) -> int:
    """Multicast address of `offset_bytes` into the symmetric buffer, or 0 when the handle has no multicast mapping."""
    try:
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
    except Exception:
        mc = 0
    return mc + offset_bytes if mc else 0


class TPShardedStream(nn.Module):
    """A LIST of R independent layers of one shape (fused QKV / gate+up projections, experts, the layers of a batch
    of models), N-sharded: this rank holds rows [start, stop) of every layer.  ``forward(xs)`` runs ONE
    ``tb200_gemv_stream_tp`` over the rank's shards of all layers, whose epilogues store every output element into
    every rank's ``[R, M, world*n_shard]`` buffer over NVLink peer memory, then ONE cross-rank barrier for the whole
    list, and returns the full ``[R, M, N]`` outputs on every rank.  Decode shapes only (M <= 32)."""

    def __init__(self, packed_shards: torch.Tensor, scale_shards, bias_shards, in_features: int, out_features: int,
                 maxval: float, group=None, max_tokens: int = 32, dtype=torch.bfloat16, multicast="auto"):
        super().__init__()
        import torch.distributed._symmetric_memory as symm_mem

        self.in_features, self.out_features, self.maxval = in_features, out_features, maxval
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_shard = shard_size(out_features, self.world)
        self.start, self.stop = shard_rows(out_features, self.world, self.rank)
        assert packed_shards.dim() == 3 and packed_shards.shape[1] == self.stop - self.start
        self.register_buffer("packed", packed_shards)
        self.register_buffer("scale", scale_shards)
        self.register_buffer("bias", bias_shards)
        R, dev = packed_shards.shape[0], packed_shards.device
        self.ld = self.world * self.n_shard
        self.max_tokens, self.dtype = max_tokens, dtype
        esz = torch.empty((), dtype=dtype).element_size()
        grp = group if group is not None else dist.group.WORLD
        self._bufs, self._hdls, self._ptrs = [], [], []
        for _ in range(2):                                   # alternate per call (see TPShardedLinear.enable_peer_gather)
            buf = symm_mem.empty((R, max_tokens, self.ld), dtype=dtype, device=dev)
            hdl = symm_mem.rendezvous(buf, grp)
            self._bufs.append(buf)
            self._hdls.append(hdl)
            self._ptrs.append((ctypes.c_void_p * self.world)(*[int(hdl.buffer_ptrs[r]) + self.rank * self.n_shard * esz
                                                               for r in _rotated(self.rank, self.world)]))
        # NVSwitch multicast mapping of the same buffers: one store reaches every rank (egress 1x instead of world x)
        self._mc = [_multicast_base(h, self.rank * self.n_shard * esz) for h in self._hdls]
        # measured (bench_tp.py): the multicast form wins on 8 GPUs (egress 1x instead of 8x: 13.7 -> 16.8 TB/s) and loses on
        # 2 / 4, where the per-rank TMA tile stores move few copies in larger pieces (prefill M=4096 on 2 GPUs: 113 vs 170 us)
        want = multicast is True or multicast == "try" or (multicast == "auto" and self.world > 4)
        self.multicast = bool(want and all(self._mc) and (self.stop - self.start) % (16 // esz) == 0)
        if multicast is True and not self.multicast:
            raise TernaryB200Error("multicast requested but the symmetric-memory handle has no multicast mapping "
                                   "(or the shard width is not a multiple of 16 bytes)")
        self._turn = 0
        self._ws = None

    def forward(self, xs: torch.Tensor, barrier: bool = True, alias_ok: bool = False) -> torch.Tensor:
        """``[R, M, N]`` gathered outputs.  They live in one of two internal symmetric buffers that peers overwrite TWO
        calls later: a copy is returned unless ``alias_ok=True`` (then the view is valid until the second following
        call; what a consumer that reads the result immediately, or a benchmark, wants)."""
        M, K = xs.shape[-2], xs.shape[-1]
        assert M <= self.max_tokens and K == self.in_features
        i = self._turn
        self._turn ^= 1
        R = self.packed.shape[0]
        need = L.lib().tb200_gemv_stream_workspace(M, K, R)
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty((need,), dtype=torch.uint8, device=xs.device)
        if self.stop > self.start:
            if self.multicast:
                ops.gemv_stream_mc(xs, self.packed, self.scale, self.bias, self.maxval, self._mc[i], self.dtype,
                                   self.ld, self.max_tokens * self.ld, workspace=self._ws)
            else:
                ops.gemv_stream_tp(xs, self.packed, self.scale, self.bias, self.maxval, self._ptrs[i], self.world, self.dtype,
                                   self.ld, self.max_tokens * self.ld, workspace=self._ws)
        if barrier:
            self._hdls[i].barrier(channel=0)             # every rank's blocks of every layer have landed in every buffer
        y = self._bufs[i][:, :M, : self.out_features]
        return y if alias_ok else y.clone()
