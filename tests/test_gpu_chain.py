"""GPU parity of tb200_chain_forward (a chain of dependent layers in one persistent launch; SURVEY.md §8f-1, the
fused MLP block) against oracle/ref_chain.py.

Layer by layer: the kernel's stored inter-layer activations h_l are compared with the oracle's (identical except
where CUDA's tanhf and numpy's differ in the last fp32 bit right at a rounding boundary of the activation dtype:
<= 1 ulp of that dtype on a vanishing fraction of elements), and every layer is then checked EXACTLY against the
oracle layer fed the kernel's own stored input (integer arithmetic is bit-exact, the epilogue rounds once)."""
import numpy as np
import pytest
import torch

from oracle import ref_chain, ref_int8
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu
DT = {torch.bfloat16: "bf16", torch.float16: "f16", torch.float32: "f32"}
EPS = {torch.bfloat16: 2.0 ** -7, torch.float16: 2.0 ** -10, torch.float32: 4e-6}


def build(dims, acts, seed, cuda, bias_on=True, scale_mode="sqrt"):
    g = torch.Generator().manual_seed(seed)
    gpu, cpu = [], []
    for i, (K, N) in enumerate(zip(dims[:-1], dims[1:])):
        W = torch.randn(N, K, generator=g).clip(-1, 1)
        scale = torch.full((N,), 1.0 / K ** 0.5) if scale_mode == "sqrt" else torch.rand(N, generator=g) * 0.2
        bias = torch.randn(N, generator=g) * 0.5 if bias_on else None
        pw = ops.quantize_pack(W.to(cuda), "ternary")
        gpu.append((pw, scale.to(cuda), None if bias is None else bias.to(cuda), 100.0, acts[i]))
        cpu.append((ref_int8.codes_of(W.numpy()), scale.numpy(), None if bias is None else bias.numpy(), 100.0, acts[i]))
    return gpu, cpu


def check_chain(plan, cpu_layers, x, out, dtype):
    dt = DT[dtype]
    M = x.shape[0]
    h_in = x.float().cpu().numpy()
    _, hs_o, _ = ref_chain.chain_forward(h_in, cpu_layers, dt)
    for l, lay in enumerate(cpu_layers):
        last = l == len(cpu_layers) - 1
        got = (out if last else plan.hidden(l, M)).float().cpu().numpy()
        want, _ = ref_chain.layer_forward(h_in, *lay, dt)            # exact: the oracle layer on the kernel's own input
        tol = np.abs(want) * EPS[dtype] + 1e-6
        bad = np.abs(got - want) > tol
        assert bad.mean() <= 2e-4, f"layer {l}: {bad.sum()} of {bad.size} outputs off by more than 1 ulp"
        assert np.all(np.abs(got - want) <= 4 * tol), f"layer {l}: max err {np.abs(got - want).max()}"
        # and against the oracle's own chain: the first layer sees identical input; later layers see inputs in which a tanh
        # boundary case may have moved a few int8 codes by one (a fraction of an output ulp each), so only closeness holds
        drift = np.abs(got - hs_o[l])
        if l == 0:
            assert np.mean(drift > tol) <= 2e-4
        assert drift.max() <= 0.02 * max(1.0, np.abs(hs_o[l]).max()), f"layer {l} drifts from the oracle chain"
        h_in = got


@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 7, 8])
@pytest.mark.parametrize("dims,acts", [
    ((256, 512, 256), ("gelu_new", None)),                      # an MLP block
    ((1024, 4096, 1024), ("gelu_new", None)),
    ((320, 200, 96, 48), ("relu", "gelu_new", None)),           # ragged N / K (K % 16 != 0 tails), three layers
    ((4096, 704, 4096), (None, None)),                          # K = 4096 -> nks 1; fewer row blocks than SMs
    ((512, 4500, 512), ("gelu_new", "relu")),                   # K = 4500 -> two k-slices, activation on the last layer
])
def test_chain_matches_oracle(cuda, M, dims, acts):
    gpu, cpu = build(dims, acts, seed=sum(dims) + M, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], torch.bfloat16)
    x = torch.randn(M, dims[0], generator=torch.Generator().manual_seed(M)).bfloat16().to(cuda)
    out = plan.forward(x)
    torch.cuda.synchronize()
    assert out.shape == (M, dims[-1]) and out.dtype == torch.bfloat16
    check_chain(plan, cpu, x, out, torch.bfloat16)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
def test_chain_dtypes(cuda, dtype):
    dims, acts = (512, 1024, 512), ("gelu_new", None)
    gpu, cpu = build(dims, acts, seed=3, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], dtype)
    x = torch.randn(4, dims[0], generator=torch.Generator().manual_seed(1)).to(dtype).to(cuda)
    out = plan.forward(x)
    check_chain(plan, cpu, x, out, dtype)
    out32 = plan.forward(x, out_dtype=torch.float32)             # fp32 output of the last layer
    assert out32.dtype == torch.float32 and torch.allclose(out32.to(dtype).float(), out.float(), rtol=2 * EPS[dtype], atol=1e-5)


def test_chain_equals_sequence_of_single_layer_calls(cuda):
    """No activation: the chain is exactly the per-layer kernels applied one after the other (bit-identical)."""
    dims = (1024, 2048, 1024, 512)
    gpu, _ = build(dims, (None, None, None), seed=5, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], torch.bfloat16)
    x = torch.randn(3, dims[0], generator=torch.Generator().manual_seed(2)).bfloat16().to(cuda)
    y = x
    for (pw, sc, bi, clip, _) in gpu:
        y = ops.gemv(y, pw, sc, bi, clip)
    assert torch.equal(plan.forward(x), y)


def test_chain_repeated_calls_graph_and_pdl(cuda):
    """The control block is cumulative across calls (generation counter): eager repeats, back-to-back programmatic
    dependent launches and CUDA-graph replays all give the same bits."""
    dims, acts = (1024, 4096, 1024), ("gelu_new", None)
    gpu, _ = build(dims, acts, seed=6, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], torch.bfloat16)
    x = torch.randn(1, dims[0], generator=torch.Generator().manual_seed(3)).bfloat16().to(cuda)
    want = plan.forward(x).clone()
    for _ in range(5):
        assert torch.equal(plan.forward(x), want)
    y = x
    outs = []
    for _ in range(6):                                           # x -> mlp -> mlp -> ... (K == N): dependent CALLS under PDL
        y = plan.forward(y, pdl=True)
        outs.append(y)
    z = x
    for i in range(6):
        z = plan.forward(z)
        assert torch.equal(z, outs[i])
    st = torch.cuda.Stream(cuda)
    out = torch.empty_like(want)
    with torch.cuda.stream(st):
        plan.forward(x, out=out)
        st.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            plan.forward(x, out=out)
        for _ in range(4):
            out.zero_()
            gr.replay()
            st.synchronize()
            assert torch.equal(out, want)


def test_chain_deep_stack_full_size(cuda):
    """16 dependent layers of the 7B-class MLP shapes (4096 -> 11008 -> 4096 x 8): every layer exact against the oracle
    layer on the kernel's own stored input."""
    dims = (4096,) + (11008, 4096) * 8
    acts = ("gelu_new", None) * 8
    gpu, cpu = build(dims, acts, seed=7, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], torch.bfloat16)
    for M in (1, 4):
        x = torch.randn(M, dims[0], generator=torch.Generator().manual_seed(M)).bfloat16().to(cuda)
        out = plan.forward(x)
        check_chain(plan, cpu, x, out, torch.bfloat16)


def test_chain_rejects_bad_arguments(cuda):
    gpu, _ = build((256, 512, 256), (None, None), seed=8, cuda=cuda)
    from ternary_vllm_b200 import TernaryB200Error
    with pytest.raises(TernaryB200Error):
        ops.ChainPlan([gpu[0], gpu[0]], 256, torch.bfloat16)                 # K of layer 1 != N of layer 0
    plan = ops.ChainPlan(gpu, 256, torch.bfloat16)
    with pytest.raises(TernaryB200Error):
        plan.forward(torch.zeros(9, 256, dtype=torch.bfloat16, device=cuda))


# ---- the module-level API: ChainedLinears / TernaryMLP (ternary_vllm_b200/blocks.py) ----
def _mlp(cuda, d, f, seed):
    import ternary_vllm_b200 as tb
    torch.manual_seed(seed)
    fc1, fc2 = tb.ScaledTernaryLinear(d, f).to(cuda), tb.ScaledTernaryLinear(f, d).to(cuda)
    fc1.scale.data.fill_(1.0 / d ** 0.5)
    fc2.scale.data.fill_(1.0 / f ** 0.5)
    return tb.TernaryMLP(fc1, fc2, "gelu_new").eval()


@pytest.mark.parametrize("M", [1, 3])
def test_mlp_module_matches_oracle_and_layerwise(cuda, M):
    d, f = 512, 1536
    mlp = _mlp(cuda, d, f, 5)
    mlp.chain_max_tokens = 8
    x = torch.randn(M, d, generator=torch.Generator().manual_seed(2)).bfloat16().to(cuda)
    with torch.no_grad():
        y = mlp(x)                                   # one persistent launch
        h = torch.nn.functional.gelu(mlp.fc1(x), approximate="tanh")
        y2 = mlp.fc2(h)                              # layer by layer through the drop-in layers
    cpu = [(ref_int8.codes_of(l.weights.detach().cpu().numpy()), l.scale.detach().cpu().numpy(), l.bias.detach().cpu().numpy(),
            l.maxval, a) for l, a in zip(mlp.layers, mlp.activations)]
    plan = mlp._plan(x.dtype, x.device)
    check_chain(plan, cpu, x, y, torch.bfloat16)
    # CUDA tanhf in the kernel vs torch's gelu: the hidden bf16 values may differ by an ulp on a few elements
    err = (y.float() - y2.float()).abs().max().item()
    assert err <= 0.02 * max(1.0, y2.float().abs().max().item())


def test_mlp_module_repacks_and_falls_back(cuda):
    d, f = 256, 512
    mlp = _mlp(cuda, d, f, 7)
    x = torch.randn(2, d, generator=torch.Generator().manual_seed(3)).bfloat16().to(cuda)
    with torch.no_grad():
        y0 = mlp(x).clone()
        mlp.fc1.weights.data[:] = torch.randn(f, d, generator=torch.Generator().manual_seed(9)).clip(-1, 1).to(cuda)   # the reference's write idiom
        y1 = mlp(x)
        assert not torch.equal(y0, y1)
        cpu = [(ref_int8.codes_of(l.weights.detach().cpu().numpy()), l.scale.detach().cpu().numpy(), l.bias.detach().cpu().numpy(),
                l.maxval, a) for l, a in zip(mlp.layers, mlp.activations)]
        check_chain(mlp._plan(x.dtype, x.device), cpu, x, y1, torch.bfloat16)
        # more tokens than the chain kernel serves: layer-by-layer path, same module
        xb = torch.randn(40, d, generator=torch.Generator().manual_seed(4)).bfloat16().to(cuda)
        yb = mlp(xb)
        assert yb.shape == (40, d)
        ref = mlp.fc2(torch.nn.functional.gelu(mlp.fc1(xb), approximate="tanh"))
        assert torch.equal(yb, ref)
    # training: autograd path (fake-quant STE), gradients reach the latent weights
    mlp.train()
    xt = torch.randn(2, d, device=cuda)
    mlp(xt).sum().backward()
    assert mlp.fc1.weights.grad is not None and mlp.fc2.weights.grad is not None


def test_fuse_mlp_blocks_in_a_model(cuda):
    """The reference's model shape: blocks with `mlp.fc1` / `mlp.activation_fn` / `mlp.fc2` (HF PhiMLP); after surgery the
    decode call of every block is one chain launch, larger calls and training keep the module's own forward."""
    import ternary_vllm_b200 as tb

    class NewGELUActivation(torch.nn.Module):
        def forward(self, x):
            return torch.nn.functional.gelu(x, approximate="tanh")

    class PhiMLP(torch.nn.Module):
        def __init__(self, d, f):
            super().__init__()
            self.activation_fn = NewGELUActivation()
            self.fc1 = torch.nn.Linear(d, f)
            self.fc2 = torch.nn.Linear(f, d)

        def forward(self, hidden_states):
            return self.fc2(self.activation_fn(self.fc1(hidden_states)))

    torch.manual_seed(3)
    model = torch.nn.Sequential(PhiMLP(256, 640), PhiMLP(256, 640)).to(cuda).eval()
    for blk in model:
        assert tb.linear_to_quantized(blk, "ternary", "none", "none")
    x = torch.randn(1, 256, device=cuda).bfloat16()
    model = model.bfloat16()
    with torch.no_grad():
        ref = model(x)                                   # layer by layer
        before = ops.launch_count()
        ref = model(x)
        per_layer = ops.launch_count() - before
        assert tb.fuse_mlp_blocks(model) == 2 and tb.fuse_mlp_blocks(model) == 0
        model(x)                                         # builds the plans
        before = ops.launch_count()
        y = model(x)
        assert ops.launch_count() - before == 2 < per_layer          # one chain launch per block
        err = (y.float() - ref.float()).abs().max().item()
        assert err <= 0.03 * max(1.0, ref.float().abs().max().item())
        xb = torch.randn(9, 256, device=cuda).bfloat16()
        assert model(xb).shape == (9, 256)               # more tokens: the module's own forward
    assert len(list(model.parameters())) == 8 + 4        # weights/bias/scale x 4 layers... parameters untouched by the surgery


@pytest.mark.parametrize("M", [1, 8])
def test_chain_full_size_mlp_block(cuda, M):
    """BASELINE-size dependent layers (the 7B-class MLP block 4096 -> 11008 -> gelu_new -> 4096, twice): every layer exactly
    against the oracle layer fed the kernel's own stored input."""
    dims, acts = (4096, 11008, 4096, 11008, 4096), ("gelu_new", None, "gelu_new", None)
    gpu, cpu = build(dims, acts, seed=77 + M, cuda=cuda)
    plan = ops.ChainPlan(gpu, dims[0], torch.bfloat16)
    x = torch.randn(M, dims[0], generator=torch.Generator().manual_seed(100 + M)).bfloat16().to(cuda)
    out = plan.forward(x)
    torch.cuda.synchronize()
    check_chain(plan, cpu, x, out, torch.bfloat16)
    out2 = plan.forward(x)                                  # second generation (the other buffer set): same bits
    assert torch.equal(out, out2)
