"""Host-side logic of the N-sharded tensor-parallel layer on CPU: world_size-2 gloo processes.
The compute of each shard is replaced by the CPU oracle (test stand-in injected through the
`matmul` hook); what is under test is the row partition, padding and the all-gather assembly."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ternary_vllm_b200.parallel import shard_rows, shard_size


def test_shard_rows_cover_and_align():
    for N in (11008, 4096, 32000, 100, 17, 1):
        for world in (1, 2, 4, 8):
            blocks = [shard_rows(N, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == N
            for (a0, b0), (a1, b1) in zip(blocks, blocks[1:]):
                assert b0 == a1 and a0 <= b0
            per = shard_size(N, world)
            assert per % 16 == 0 and all(b - a <= per for a, b in blocks)
    assert shard_rows(11008, 8, 3) == (4128, 5504)           # 1376 rows per rank
    assert shard_rows(32000, 8, 7) == (28000, 32000)
    assert shard_rows(17, 4, 2) == (17, 17)                  # empty trailing shard


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, N, K, M, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ternary_vllm_b200 as tb
        from ternary_vllm_b200.parallel import TPShardedLinear
        from oracle import ref_modules as om

        def cpu_matmul(x, w_rows, scale, bias, clip):       # stand-in for ops.ternary_matmul (tests only)
            return om.packed_ternary_matmul(x, om.ternary_codes(w_rows), scale, 0 if bias is None else bias, clip)

        torch.manual_seed(0)                                 # same full layer on every rank
        layer = tb.ScaledTernaryLinear(K, N)
        layer.scale.data = torch.rand(N) * 0.1
        x = torch.randn(2, M, K)
        tp = TPShardedLinear.from_layer(layer, matmul=cpu_matmul)
        y = tp(x)
        want = om.packed_ternary_matmul(x.reshape(-1, K), om.ternary_codes(layer.weights.detach()), layer.scale.detach(),
                                        layer.bias.detach(), layer.maxval).reshape(2, M, N)
        ok = torch.allclose(y, want, rtol=1e-5, atol=1e-5) and y.shape == (2, M, N)
        ret[rank] = (ok, tp.start, tp.stop)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("N,K,M", [(100, 32, 3), (48, 64, 1), (17, 16, 2), (16, 16, 2)])      # (16, .): rank 1 owns no rows
def test_tp_all_gather_world2(N, K, M):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), N, K, M, ret), nprocs=world, join=True)
    assert len(ret) == world and all(v[0] for v in ret.values()), dict(ret)
    assert ret[0][1] == 0 and ret[0][2] == ret[1][1] and ret[1][2] == N


def test_peer_gather_path_choice_is_rank_uniform():
    """the fused-gather path must not depend on whether THIS rank's shard is empty (an NCCL all-gather on
    one rank against a symmetric-memory barrier on the others hangs)."""
    import inspect
    from ternary_vllm_b200.parallel import TPShardedLinear
    src = inspect.getsource(TPShardedLinear.uses_peer_gather)
    assert "self.stop" not in src and "self.start" not in src
