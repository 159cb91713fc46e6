"""GPU parity of the drop-in layer classes and model surgery (mirrors how train.py / convert.py use them)."""
import numpy as np
import pytest
import torch
from torch import nn

import ternary_vllm_b200 as tb
from oracle import ref_int8, ref_modules as om
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu


def oracle_layer(layer, x, mode):
    q, s, _ = ref_int8.quant_act(x.float().cpu().numpy())
    codes = ref_int8.codes_of(layer.weights.detach().cpu().numpy(), mode)
    scale = layer.scale.detach().cpu().numpy() if hasattr(layer, "scale") else None
    bias = None if layer.bias is None else layer.bias.detach().cpu().numpy()
    return ref_int8.fused_out(ref_int8.int_acc(q, codes), s, scale, bias, layer.maxval)


@pytest.mark.parametrize("cls,mode", [(tb.ScaledTernaryLinear, "ternary"), (tb.ScaledBinaryLinear, "binary"),
                                      (tb.ScaledBinary01Linear, "binary01"), (tb.SmoothlyBinarizedLinear, "binary")])
@pytest.mark.parametrize("shape", [(1, 320), (2, 5, 320), (3, 40, 320)])
def test_layer_forward(cuda, cls, mode, shape):
    torch.manual_seed(1)
    layer = cls(320, 200).to(cuda).eval()
    if hasattr(layer, "scale"):
        layer.scale.data.fill_(0.05)
    x = torch.randn(*shape, device=cuda, dtype=torch.bfloat16)
    with torch.no_grad():
        y = layer(x)
    assert y.shape == (*shape[:-1], 200) and y.dtype == torch.bfloat16
    want = oracle_layer(layer, x.reshape(-1, 320), mode)
    got = y.float().cpu().numpy().reshape(-1, 200)
    assert np.all(np.abs(got - want) <= 2.0 ** -7 * np.abs(want) + 1e-5)


def test_layer_matches_reference_golden(cuda, golden):
    W, x = torch.from_numpy(golden["small_W"]), torch.from_numpy(golden["small_x"])
    layer = tb.ScaledTernaryLinear(W.shape[1], W.shape[0]).to(cuda)
    layer.load_state_dict({"weights": W, "scale": torch.from_numpy(golden["small_scale"]), "bias": torch.from_numpy(golden["small_bias"])})
    ref = golden["small_y_ternary"]
    rms = np.sqrt((ref ** 2).mean())
    layer.act_quant = False                                  # float activations: the reference's exact semantics
    with torch.no_grad():
        y = layer(x.to(cuda)).cpu().numpy()
    assert np.all(np.abs(y - ref) <= 1e-4 * np.maximum(np.abs(ref), rms))
    layer.act_quant = True                                   # int8 activations: error is bounded by the quantisation step
    with torch.no_grad():
        y8 = layer(x.to(cuda)).cpu().numpy()
    assert np.abs(y8 - ref).max() / rms < 0.05
    # training formulation (autograd on) reproduces the reference bit-for-bit up to fp32 rounding
    y_t = layer(x.to(cuda))
    assert np.allclose(y_t.detach().cpu().numpy(), ref, rtol=1e-5, atol=1e-5)
    y_t.sum().backward()
    assert layer.weights.grad is not None and layer.scale.grad is not None and layer.bias.grad is not None
    s = tb.SmoothlyBinarizedLinear(W.shape[1], W.shape[0]).to(cuda)
    s.load_state_dict({"weights": W, "bias": torch.from_numpy(golden["small_bias"])})
    s.train()
    assert np.allclose(s(x.to(cuda)).detach().cpu().numpy(), golden["small_y_smooth_train"], rtol=1e-5, atol=1e-5)


def test_repack_on_weight_update_and_frozen(cuda):
    torch.manual_seed(2)
    layer = tb.ScaledTernaryLinear(128, 64).to(cuda)
    x = torch.randn(2, 128, device=cuda)
    with torch.no_grad():
        y0 = layer(x).clone()
        layer.weights.mul_(-1.0)                 # in-place op bumps the version -> repack
        y1 = layer(x)
    assert not torch.equal(y0, y1)
    assert np.array_equal(layer.codes().cpu().numpy(), ref_int8.codes_of(layer.weights.detach().cpu().numpy()))
    layer.weights.requires_grad_(False)          # frozen weights are fine here
    with torch.no_grad():
        assert torch.equal(layer(x), y1)
    layer.freeze()
    assert torch.equal(layer(x), y1)             # frozen: packed path even with autograd enabled
    # input must not be mutated
    x2 = x.clone()
    layer(x2)
    assert torch.equal(x, x2)


CASES = [("ternary", "none", "uniform"), ("ternary", "none", "none"), ("binary", "none", "uniform"),
         ("binary", "std", "independent"), ("binary", "natural", "independent"), ("binary01", "none", "independent"),
         ("smoothBinary", "none", "uniform")]


@pytest.mark.parametrize("qn,sc,ns", CASES)
def test_linear_to_quantized_golden(cuda, golden, qn, sc, ns):
    holder = nn.Sequential(nn.Linear(40, 24)).to(cuda)
    holder[0].weight.data = torch.from_numpy(golden["l2q_W"]).to(cuda)
    holder[0].bias.data = torch.from_numpy(golden["l2q_b"]).to(cuda)
    assert tb.linear_to_quantized(holder, quantization=qn, scaling=sc, neuron_scale=ns) is True
    new = holder[0]
    cls = {"ternary": tb.ScaledTernaryLinear, "binary": tb.ScaledBinaryLinear, "binary01": tb.ScaledBinary01Linear,
           "smoothBinary": tb.SmoothlyBinarizedLinear}[qn]
    assert type(new) is cls and new.weights.device.type == "cuda"
    key = f"l2q_{qn}_{sc}_{ns}"
    assert np.allclose(new.weights.detach().cpu().numpy(), golden[key + "_latent"], rtol=2e-6, atol=1e-7)
    if hasattr(new, "scale"):
        assert np.allclose(new.scale.detach().cpu().numpy(), golden[key + "_scale"], rtol=2e-6)
    assert np.array_equal(new.bias.detach().cpu().numpy(), golden[key + "_bias"])
    with torch.no_grad():
        new(torch.randn(3, 40, device=cuda))


def test_linear_to_quantized_bf16_source(cuda, golden):
    holder = nn.Sequential(nn.Linear(40, 24)).to(cuda).to(torch.bfloat16)
    holder[0].weight.data = torch.from_numpy(golden["l2q_W"]).to(cuda).to(torch.bfloat16)
    tb.linear_to_quantized(holder, quantization="ternary", scaling="none", neuron_scale="none")
    assert np.allclose(holder[0].scale.detach().float().cpu().numpy(), golden["l2q_bf16_ternary_scale"], rtol=1e-5)
    assert np.allclose(holder[0].weights.detach().float().cpu().numpy(), golden["l2q_bf16_ternary_latent"], rtol=1e-5, atol=1e-6)


def test_clip_binary_weights_golden(cuda, golden):
    m = nn.Sequential(tb.ScaledTernaryLinear(20, 10), tb.SmoothlyBinarizedLinear(10, 6)).to(cuda)
    m[0].weights.data = torch.from_numpy(golden["clip_in0"]).to(cuda)
    m[1].weights.data = torch.from_numpy(golden["clip_in1"]).to(cuda)
    with torch.no_grad():
        m[0](torch.randn(1, 20, device=cuda))          # builds the packed cache, which the clip must invalidate
    tb.clip_binary_weights(m)
    assert np.array_equal(m[0].weights.detach().cpu().numpy(), golden["clip_out0"])
    assert np.array_equal(m[1].weights.detach().cpu().numpy(), golden["clip_out1"])
    assert np.array_equal(m[0].codes().cpu().numpy(), ref_int8.codes_of(golden["clip_out0"]))


class _Mixer(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.Wqkv, self.out_proj, self.layer_idx = nn.Linear(d, 3 * d), nn.Linear(d, d), -1


class _MLP(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.fc1, self.fc2 = nn.Linear(d, 4 * d), nn.Linear(4 * d, d)


class _Block(nn.Module):
    def __init__(self, d):
        super().__init__()
        self.ln, self.mixer, self.mlp = nn.LayerNorm(d), _Mixer(d), _MLP(d)


class _FakeMoondream(nn.Module):
    """Module tree with the names quantize_moondream walks (binary/utils.py:8-33, conversion.py:133-136)."""

    def __init__(self, d=64, n=4):
        super().__init__()
        self.text_model = nn.Module()
        self.text_model.transformer = nn.Module()
        self.text_model.transformer.embd = nn.Embedding(100, d)
        self.text_model.transformer.h = nn.ModuleList([_Block(d) for _ in range(n)])
        self.text_model.lm_head = nn.Module()
        self.text_model.lm_head.ln = nn.LayerNorm(d)
        self.text_model.lm_head.linear = nn.Linear(d, 100)


def test_quantize_moondream(cuda, capsys):
    torch.manual_seed(3)
    model = _FakeMoondream().to(cuda)
    acc = tb.quantize_moondream(model, start_skip=1, last_skip=1, quantization="ternary", scaling="none",
                                neuron_scale="none", remove_blocks=[2])
    h = model.text_model.transformer.h
    assert len(h) == 3 and [b.mixer.layer_idx for b in h] == [0, 1, 2]
    assert isinstance(h[0].mlp.fc1, nn.Linear)                       # first block skipped
    assert isinstance(h[1].mlp.fc1, tb.ScaledTernaryLinear) and isinstance(h[1].mixer.Wqkv, tb.ScaledTernaryLinear)
    assert isinstance(h[2].mlp.fc1, nn.Linear)                       # last block skipped (was index 3)
    assert isinstance(model.text_model.lm_head.linear, nn.Linear)    # lm_head never converted
    assert acc["n_ternary_modules"] == 4 and "estimated quantized size" in capsys.readouterr().out
    with torch.no_grad():
        y = h[1].mlp.fc2(torch.nn.functional.gelu(h[1].mlp.fc1(torch.randn(5, 64, device=cuda))))
    assert y.shape == (5, 64) and torch.isfinite(y).all()


# ---- the packed-weight cache under the reference's own write idioms -----------------------
@pytest.mark.parametrize("idiom", ["data_slice", "data_copy", "data_assign", "inplace", "load_state_dict", "ref_clip"])
def test_packed_cache_follows_reference_write_idioms(cuda, idiom):
    """forward -> mutate the latent weights the way the reference does -> forward must use the NEW codes.
    `p.data[:] = v` is convert.py:78 (checkpoint load by parameter name); `weights.data = ...` is
    binary/modules.py:226,261,292 and binary/utils.py:77.  None of the .data forms bumps weights._version."""
    torch.manual_seed(5)
    layer = tb.ScaledTernaryLinear(256, 96).to(cuda)
    x = torch.randn(3, 256, device=cuda, dtype=torch.bfloat16)
    with torch.no_grad():
        y0 = layer(x).clone()
    w2 = torch.randn(96, 256, device=cuda).clip(-1, 1)
    if idiom == "data_slice":
        for name, p in layer.named_parameters():              # convert.py:76-78
            if name == "weights":
                p.data[:] = w2
    elif idiom == "data_copy":
        layer.weights.data.copy_(w2)
    elif idiom == "data_assign":
        layer.weights.data = w2.clone()
    elif idiom == "inplace":
        with torch.no_grad():
            layer.weights.copy_(w2)
    elif idiom == "load_state_dict":
        sd = layer.state_dict()
        sd["weights"] = w2.clone()
        layer.load_state_dict(sd)
    else:                                                      # the reference's clip_binary_weights body verbatim
        layer.weights.data = (layer.weights.data * 3).clip(-1, 1)
        w2 = layer.weights.detach().clone()
    with torch.no_grad():
        y1 = layer(x)
    assert np.array_equal(layer.codes().cpu().numpy(), ref_int8.codes_of(w2.cpu().numpy()))
    want = oracle_layer(layer, x, "ternary")
    assert np.all(np.abs(y1.float().cpu().numpy() - want) <= 2.0 ** -7 * np.abs(want) + 1e-5)
    assert not torch.equal(y0, y1)


def test_freeze_state_dict_round_trip_and_release(cuda):
    torch.manual_seed(6)
    layer = tb.ScaledTernaryLinear(128, 64).to(cuda)
    x = torch.randn(2, 128, device=cuda, dtype=torch.bfloat16)
    assert sorted(layer.state_dict()) == ["bias", "scale", "weights"]          # the reference's checkpoint keys only
    with torch.no_grad():
        y = layer(x)
    layer.freeze(release_latent=True)
    sd = layer.state_dict()
    assert "_packed" in sd and sd["weights"].numel() == 0
    assert torch.equal(layer(x), y)
    with pytest.raises(tb.TernaryB200Error):
        layer.invalidate_packed()
    fresh = tb.ScaledTernaryLinear(128, 64).to(cuda)
    fresh.load_state_dict(sd)
    assert torch.equal(fresh(x), y) and fresh._frozen and fresh._packed.device == x.device


def test_torch_compile_keeps_the_layer_as_one_op(cuda):
    """the ops are registered with torch.library, so torch.compile captures the layer without a graph
    break (fullgraph=True would raise otherwise)."""
    torch.manual_seed(8)
    layer = tb.ScaledTernaryLinear(256, 128).to(cuda).eval()
    layer.freeze()
    x = torch.randn(4, 256, device=cuda, dtype=torch.bfloat16)
    want = layer(x)
    fn = torch.compile(lambda t: layer(t) * 2.0, backend="aot_eager", fullgraph=True)
    got = fn(x)
    assert torch.equal(got, want * 2.0)
    y = torch.ops.ternary_b200.ternary_matmul(x, layer.packed_weights(), layer.scale.detach(), layer.bias.detach(),
                                              100.0, torch.bfloat16)
    assert torch.equal(y, want)
    torch.library.opcheck(torch.ops.ternary_b200.ternary_matmul.default,
                          (x, layer.packed_weights(), layer.scale.detach(), layer.bias.detach(), 100.0, torch.bfloat16),
                          test_utils=("test_schema", "test_faketensor"))


def test_two_host_threads_two_streams(cuda):
    """The C ABI is called from two host threads on two streams at once (every kernel family, first use included:
    the per-device dynamic-shared-memory opt-in and tensor-map caches are shared state)."""
    import threading
    torch.manual_seed(9)
    layers = [tb.ScaledTernaryLinear(1024, 512).to(cuda).eval().freeze() for _ in range(2)]
    xs = {M: torch.randn(M, 1024, device=cuda, dtype=torch.bfloat16) for M in (1, 12, 200)}
    with torch.no_grad():
        want = [{M: l(x).clone() for M, x in xs.items()} for l in layers]
    torch.cuda.synchronize()
    errs = []

    def work(i):
        try:
            st = torch.cuda.Stream(cuda)
            with torch.cuda.stream(st), torch.no_grad():
                for _ in range(20):
                    for M, x in xs.items():
                        x.record_stream(st)
                        if not torch.equal(layers[i](x), want[i][M]):
                            errs.append((i, M))
            st.synchronize()
        except Exception as e:      # noqa: BLE001
            errs.append(repr(e))

    th = [threading.Thread(target=work, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


def test_second_device_if_present(cuda):
    """Per-device state (SM count, dynamic shared-memory opt-in): a layer on cuda:1 after one on cuda:0."""
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU")
    torch.manual_seed(10)
    a = tb.ScaledTernaryLinear(4096, 512).to("cuda:0").eval()
    b = copy_to = tb.ScaledTernaryLinear(4096, 512).to("cuda:1").eval()
    b.load_state_dict(a.state_dict())
    for M in (1, 12, 300):
        x = torch.randn(M, 4096, dtype=torch.bfloat16)
        with torch.no_grad():
            ya, yb = a(x.to("cuda:0")), b(x.to("cuda:1"))
        assert torch.equal(ya.cpu(), yb.cpu())
    with pytest.raises(tb.TernaryB200Error):
        a.freeze()
        a(torch.randn(1, 4096, dtype=torch.bfloat16, device="cuda:1"))


# ---- checkpoint path: load by parameter name straight into packed form; export of formats A / B / C ----
def _tiny_model(dev):
    import ternary_vllm_b200 as tb
    torch.manual_seed(4)
    m = torch.nn.Sequential()
    m.add_module("ln", torch.nn.LayerNorm(96))
    m.add_module("fc1", tb.ScaledTernaryLinear(96, 200))
    m.add_module("fc2", tb.ScaledBinaryLinear(200, 64))
    m.add_module("head", torch.nn.Linear(64, 10))
    return m.to(dev)


def test_load_safetensors_packed_matches_reference_loading(cuda, tmp_path):
    from safetensors.torch import save_file

    from ternary_vllm_b200.checkpoint import load_safetensors_packed
    src = _tiny_model("cpu")
    for p in src.parameters():
        p.data = torch.randn_like(p) * 0.7
    path = tmp_path / "model.safetensors"
    save_file({k: v.detach().clone() for k, v in src.named_parameters()}, str(path))
    # the reference's way (convert.py:72-79): p.data[:] = v on a model that keeps its latent weights
    ref = _tiny_model(cuda)
    from safetensors import safe_open
    with safe_open(str(path), "pt") as f:
        for name, p in ref.named_parameters():
            p.data[:] = f.get_tensor(name)
    # packed loading: no latent weights stay resident
    m = _tiny_model(cuda)
    stats = load_safetensors_packed(m, path)
    assert stats["packed_layers"] == 2 and stats["copied"] == 2 + 2 + 2 + 2 and not stats["missing"]
    assert m.fc1.weights.numel() == 0 and m.fc2.weights.numel() == 0 and m.fc1._frozen and m.fc2._frozen
    assert torch.equal(m.fc1.packed_weights(), ref.fc1.packed_weights()) and torch.equal(m.fc2.packed_weights(), ref.fc2.packed_weights())
    x = torch.randn(5, 96, device=cuda)
    with torch.no_grad():
        assert torch.equal(m(x), ref(x))
    # the frozen model round-trips through state_dict (packed buffers are persistent once frozen)
    m2 = _tiny_model(cuda)
    m2.load_state_dict(m.state_dict())
    with torch.no_grad():
        assert torch.equal(m2(x), ref(x))
    # keep the latent weights instead
    m3 = _tiny_model(cuda)
    load_safetensors_packed(m3, path, release_latent=False)
    assert torch.equal(m3.fc1.weights.detach().cpu(), src.fc1.weights.detach())
    with torch.no_grad():
        assert torch.equal(m3(x), ref(x))
    with pytest.raises(KeyError):
        save_file({"ln.weight": torch.ones(96)}, str(tmp_path / "partial.safetensors"))
        load_safetensors_packed(_tiny_model(cuda), tmp_path / "partial.safetensors")


@pytest.mark.parametrize("fmt", ["A", "B", "C"])
def test_export_packed_formats(cuda, fmt):
    from oracle import ref_packing as rp
    from ternary_vllm_b200.checkpoint import export_packed
    m = _tiny_model(cuda)
    out = export_packed(m, fmt)
    assert set(out) == {"fc1", "fc2"}
    for name, layer in (("fc1", m.fc1), ("fc2", m.fc2)):
        codes = layer.codes().cpu().numpy()
        want = {"A": lambda c: rp.pack_A(c.reshape(-1)), "B": lambda c: rp.pack_B(c.reshape(-1)), "C": rp.pack_C}[fmt](codes)
        got = out[name]["packed"].numpy()
        assert got.dtype == np.uint8 and np.array_equal(got.reshape(-1), np.asarray(want).reshape(-1))
        assert out[name]["clip_val"] == layer.maxval and out[name]["shape"] == (layer.out_features, layer.in_features)


# ---- fused training helpers: compute_and_set_biobjective_gradient on the GPU ----
def test_biobjective_gradient_fused_golden(cuda, golden):
    from ternary_vllm_b200 import modules as tbm
    torch.manual_seed(7)
    net = nn.Sequential(tb.SmoothlyBinarizedLinear(12, 8), nn.Linear(8, 4)).to(cuda)
    net[0].weights.data = torch.from_numpy(golden["bio_w0"]).clone().to(cuda)
    g_in = torch.from_numpy(golden["bio_g_in"]).to(cuda)
    off = 0
    for p in net.parameters():
        p.grad = g_in[off:off + p.numel()].reshape(p.shape).clone()
        off += p.numel()
    before = ops.launch_count()
    assert tbm._biobjective_fused(net)                       # the CUDA kernels, not the torch formulation
    assert ops.launch_count() - before == 4 + 1 + 4          # 4 gradient norms, 1 sign count, 4 in-place updates
    got = torch.cat([p.grad.flatten() for p in net.parameters()]).cpu().numpy()
    assert np.allclose(got, golden["bio_g_out"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("zero_g", [False, True])
def test_biobjective_gradient_fused_matches_torch_formulation(cuda, zero_g):
    """Large tensors (vector path + tail, several CTAs), incl. the norm_g == 0 branch, against the oracle restatement."""
    torch.manual_seed(11)
    net = nn.Sequential(tb.SmoothlyBinarizedLinear(1000, 777), nn.LayerNorm(777), tb.SmoothlyBinarizedLinear(777, 33, bias=False)).to(cuda)
    for p in net.parameters():
        p.grad = torch.zeros_like(p) if zero_g else torch.randn_like(p) * 0.01
    net[0].weights.data[0, :5] = 0.0                          # sign(0) = 0 entries
    smooth_w = [net[0].weights.detach().cpu(), net[2].weights.detach().cpu()]
    smooth_g = [net[0].weights.grad.cpu().clone(), net[2].weights.grad.cpu().clone()]
    others = [p.grad.cpu().clone() for n, p in net.named_parameters() if not n.endswith("weights")]
    want_s, want_o = om.biobjective_gradients(smooth_w, smooth_g, others)
    tb.compute_and_set_biobjective_gradient(net)
    got_s = [net[0].weights.grad.cpu(), net[2].weights.grad.cpu()]
    got_o = [p.grad.cpu() for n, p in net.named_parameters() if not n.endswith("weights")]
    for a, b in zip(got_s + got_o, list(want_s) + list(want_o)):
        assert torch.allclose(a, b, rtol=2e-5, atol=1e-9), (a - b).abs().max()
