"""Host-side logic that needs no GPU: module surface, STE values, state-dict contract, errors."""
import numpy as np
import pytest
import torch
from torch import nn

import ternary_vllm_b200 as tb
from oracle import ref_modules as om


def test_constructor_and_state_dict_contract():
    m = tb.ScaledTernaryLinear(12, 7, bias=True, max_val=-50.0)
    assert m.maxval == 50.0
    assert sorted(m.state_dict()) == ["bias", "scale", "weights"]       # names are contract (convert.py:76-78)
    assert m.weights.shape == (7, 12) and m.scale.shape == (7,) and torch.all(m.scale == 1)
    assert tb.ScaledTernaryLinear(4, 4, bias=False).bias is None
    s = tb.SmoothlyBinarizedLinear(4, 3)
    assert sorted(s.state_dict()) == ["bias", "weights"] and not hasattr(s, "scale")
    for cls in (tb.ScaledBinaryLinear, tb.ScaledBinary01Linear, tb.ScaledTernaryLinear, tb.SmoothlyBinarizedLinear):
        assert isinstance(cls(3, 2), nn.Module)


def test_ste_values_and_gradients(golden):
    W = torch.from_numpy(golden["small_W"]).clone().requires_grad_(True)
    assert np.array_equal(tb.tri_step_ste(W).detach().numpy(), golden["small_q_ternary"])
    assert np.array_equal(tb.sign_ste(W).detach().numpy(), golden["small_q_sign"])
    assert np.array_equal(tb.step_ste(W).detach().numpy(), golden["small_q_step"])
    assert np.array_equal(tb.sign_ste_sat(W).detach().numpy(), golden["small_q_sign_sat"])
    assert np.array_equal(tb.step_ste_sat(W).detach().numpy(), golden["small_q_step_sat"])
    tb.tri_step_ste(W).sum().backward()
    assert torch.all(W.grad == 1)                      # straight-through
    # frozen weights work here (the reference raises UnboundLocalError)
    Wf = W.detach()
    assert torch.equal(tb.tri_step_ste(Wf), om.ternary_codes(Wf).to(Wf.dtype))


def test_layers_refuse_cpu_tensors():
    for cls in (tb.ScaledTernaryLinear, tb.ScaledBinaryLinear, tb.ScaledBinary01Linear, tb.SmoothlyBinarizedLinear):
        with pytest.raises(tb.TernaryB200Error):
            cls(8, 4)(torch.randn(2, 8))          # training mode / autograd on: still no CPU path


def test_linear_to_quantized_errors_and_noop():
    with pytest.raises(ValueError):
        tb.linear_to_quantized(nn.Sequential(nn.Linear(4, 4)), quantization="binary", neuron_scale="bogus")
    with pytest.raises(ValueError):
        tb.linear_to_quantized(nn.Sequential(nn.Linear(4, 4)), quantization="binary", scaling="bogus", neuron_scale="uniform")
    assert tb.linear_to_quantized(nn.Sequential(nn.ReLU()), quantization="ternary") is False
    assert tb.linear_to_quantized(nn.Sequential(nn.Linear(4, 4)), quantization="nonsense") is False


def test_size_accounting():
    net = nn.Sequential(tb.ScaledTernaryLinear(10, 10), tb.ScaledBinaryLinear(10, 10), nn.Linear(10, 10), nn.LayerNorm(10))
    acc = tb.quantized_size_bits(net)
    want = om.module_bits("ternary", 100, 10) + om.module_bits("binary", 100, 10) + om.module_bits("linear", 100, 10) + 20 * 16
    assert acc["n_bits"] == want and acc["n_ternary_modules"] == 1 and acc["n_binary_modules"] == 1


def test_biobjective_gradient_golden(golden):
    torch.manual_seed(7)
    net = nn.Sequential(tb.SmoothlyBinarizedLinear(12, 8), nn.Linear(8, 4))
    net[0].weights.data = torch.from_numpy(golden["bio_w0"]).clone()
    g_in = torch.from_numpy(golden["bio_g_in"])
    off = 0
    for p in net.parameters():
        p.grad = g_in[off:off + p.numel()].reshape(p.shape).clone()
        off += p.numel()
    tb.compute_and_set_biobjective_gradient(net)
    got = torch.cat([p.grad.flatten() for p in net.parameters()])
    assert np.allclose(got.numpy(), golden["bio_g_out"], rtol=1e-5, atol=1e-6)
    n, c = tb.compute_smooth_binary_stats(net, eps=0.2)
    assert [float(n), float(c)] == golden["bio_stats"].tolist()


def test_tp_peer_order_is_a_rotation():
    """Every rank stores to all ranks exactly once, and at step i the ranks address pairwise different receivers."""
    from ternary_vllm_b200.parallel import _rotated
    for world in (1, 2, 4, 8):
        orders = [_rotated(r, world) for r in range(world)]
        for r, o in enumerate(orders):
            assert sorted(o) == list(range(world)) and o[-1] == r
        for i in range(world):
            assert len({o[i] for o in orders}) == world
