"""Generate tests/golden/*.npz by running the LIVE reference (/root/reference) in the build
container.  The GPU box has no /root/reference, so the outputs are committed as fixtures.

    python oracle/make_golden.py            # rewrites tests/golden/

Everything here calls the reference's own functions; nothing is re-implemented.  Inputs are
either stored in the fixture (small cases) or regenerated from a torch CPU seed (large case:
torch's CPU generator is deterministic for a fixed torch version, recorded in the fixture).
"""
import ast
import os
import sys

import numpy as np
import torch
from torch import nn

REF = "/root/reference"
sys.path.insert(0, REF)
import binary.modules as rm  # noqa: E402
import binary.utils as ru  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def ref_forward(cls, W, scale, bias, x, maxval=100.0):
    m = cls(W.shape[1], W.shape[0], bias is not None, max_val=maxval)
    m.weights.data = W.clone()
    if hasattr(m, "scale") and scale is not None:
        m.scale.data = scale.clone()
    if bias is not None:
        m.bias.data = bias.clone()
    with torch.no_grad():            # weights keep requires_grad=True
        return m(x.clone())


def main():
    g = {}
    torch.manual_seed(1234)
    # ---- small forward cases, inputs stored ------------------------------------------------
    for tag, (N, K, M) in {"small": (48, 100, 3), "k4": (16, 64, 1), "ragged": (37, 259, 5)}.items():
        W = torch.randn(N, K).clip(-1, 1)
        W[0, :4] = torch.tensor([0.5, -0.5, 0.49999, -0.49999])     # decision boundaries
        W[1, :2] = torch.tensor([0.0, -0.0])
        x = torch.randn(M, K)
        scale = torch.rand(N) * 0.2 + 0.01
        bias = torch.randn(N)
        g[f"{tag}_W"], g[f"{tag}_x"], g[f"{tag}_scale"], g[f"{tag}_bias"] = W.numpy(), x.numpy(), scale.numpy(), bias.numpy()
        g[f"{tag}_y_ternary"] = ref_forward(rm.ScaledTernaryLinear, W, scale, bias, x).numpy()
        g[f"{tag}_y_binary"] = ref_forward(rm.ScaledBinaryLinear, W, scale, bias, x).numpy()
        g[f"{tag}_y_binary01"] = ref_forward(rm.ScaledBinary01Linear, W, scale, bias, x).numpy()
        g[f"{tag}_y_ternary_nobias"] = ref_forward(rm.ScaledTernaryLinear, W, scale, None, x).numpy()
        sm = rm.SmoothlyBinarizedLinear(K, N, True)
        sm.weights.data, sm.bias.data = W.clone(), bias.clone()
        sm.eval()
        with torch.no_grad():
            g[f"{tag}_y_smooth_eval"] = sm(x.clone()).numpy()
        sm.train()
        with torch.no_grad():
            g[f"{tag}_y_smooth_train"] = sm(x.clone()).numpy()
        # clip path: scale = 1 makes |y| exceed maxval for some outputs
        g[f"{tag}_y_ternary_clip"] = ref_forward(rm.ScaledTernaryLinear, W, torch.ones(N) * 4.0, bias, x, maxval=10.0).numpy()
        # quantiser values (through the STE functions with requires_grad=True)
        Wg = W.clone().requires_grad_(True)
        with torch.no_grad():
            g[f"{tag}_q_ternary"] = rm.tri_step_ste(Wg).numpy()
            g[f"{tag}_q_sign"] = rm.sign_ste(Wg).numpy()
            g[f"{tag}_q_step"] = rm.step_ste(Wg).numpy()
            g[f"{tag}_q_sign_sat"] = rm.sign_ste_sat(Wg).numpy()
            g[f"{tag}_q_step_sat"] = rm.step_ste_sat(Wg).numpy()

    # ---- BASELINE config 0: single ternary Linear 2048x2048, batch 1, fp32 (inputs from seed) ---
    torch.manual_seed(1234)
    W = torch.randn(2048, 2048).clip(-1, 1)
    x = torch.randn(1, 2048)
    bias = torch.randn(2048)
    scale = torch.full((2048,), 1.0 / 2048 ** 0.5)
    g["cfg0_y"] = ref_forward(rm.ScaledTernaryLinear, W, scale, bias, x).numpy()
    g["cfg0_y_scale1"] = ref_forward(rm.ScaledTernaryLinear, W, torch.ones(2048), bias, x).numpy()
    codes = (W >= 0.5).to(torch.int8) - (W <= -0.5).to(torch.int8)
    g["cfg0_code_hist"] = np.array([(codes == v).sum().item() for v in (-1, 0, 1)])
    g["cfg0_W_checksum"] = np.array([float(W.double().sum()), float(W.double().abs().sum())])
    g["torch_version"] = np.array(torch.__version__)

    # ---- linear_to_quantized: every mode, inputs stored -----------------------------------
    torch.manual_seed(4321)
    lin = nn.Linear(40, 24)
    with torch.no_grad():
        lin.weight.mul_(3.0)
    g["l2q_W"], g["l2q_b"] = lin.weight.detach().numpy().copy(), lin.bias.detach().numpy().copy()
    cases = [("ternary", "none", "uniform"), ("ternary", "none", "none"), ("ternary", "none", "independent"),
             ("binary", "none", "uniform"), ("binary", "none", "independent"), ("binary", "std", "independent"),
             ("binary", "natural", "independent"), ("binary", 0.5, "uniform"),
             ("binary01", "none", "uniform"), ("binary01", "none", "independent"), ("binary01", "std", "uniform"),
             ("smoothBinary", "none", "uniform")]
    for qn, sc, ns in cases:
        holder = nn.Sequential(nn.Linear(40, 24))
        holder[0].weight.data = lin.weight.data.clone()
        holder[0].bias.data = lin.bias.data.clone()
        assert rm.linear_to_quantized(holder, quantization=qn, scaling=sc, neuron_scale=ns)
        new = holder[0]
        key = f"l2q_{qn}_{sc}_{ns}"
        g[key + "_latent"] = new.weights.detach().numpy().copy()
        if hasattr(new, "scale"):
            g[key + "_scale"] = new.scale.detach().numpy().copy()
        g[key + "_bias"] = new.bias.detach().numpy().copy()
    # bf16 source weights (the reference is normally applied to an fp16/bf16 HF model)
    holder = nn.Sequential(nn.Linear(40, 24)).to(torch.bfloat16)
    holder[0].weight.data = lin.weight.data.to(torch.bfloat16)
    holder[0].bias.data = lin.bias.data.to(torch.bfloat16)
    rm.linear_to_quantized(holder, quantization="ternary", scaling="none", neuron_scale="none")
    g["l2q_bf16_ternary_latent"] = holder[0].weights.detach().float().numpy().copy()
    g["l2q_bf16_ternary_scale"] = holder[0].scale.detach().float().numpy().copy()

    # ---- clip_binary_weights ----------------------------------------------------------------
    torch.manual_seed(99)
    m = nn.Sequential(rm.ScaledTernaryLinear(20, 10), rm.SmoothlyBinarizedLinear(10, 6))
    m[0].weights.data = torch.randn(10, 20) * 2
    m[1].weights.data = torch.randn(6, 10) * 2
    g["clip_in0"], g["clip_in1"] = m[0].weights.detach().numpy().copy(), m[1].weights.detach().numpy().copy()
    ru.clip_binary_weights(m)
    g["clip_out0"], g["clip_out1"] = m[0].weights.detach().numpy().copy(), m[1].weights.detach().numpy().copy()

    # ---- bi-objective gradient --------------------------------------------------------------
    torch.manual_seed(7)
    net = nn.Sequential(rm.SmoothlyBinarizedLinear(12, 8), nn.Linear(8, 4))
    for p in net.parameters():
        p.grad = torch.randn_like(p)
    g["bio_w0"] = net[0].weights.detach().numpy().copy()
    g["bio_g_in"] = np.concatenate([p.grad.flatten().numpy() for p in net.parameters()])
    rm.compute_and_set_biobjective_gradient(net)
    g["bio_g_out"] = np.concatenate([p.grad.flatten().numpy() for p in net.parameters()])
    g["bio_stats"] = np.array([float(v) for v in rm.compute_smooth_binary_stats(net, eps=0.2)])

    # ---- ternary_mm of matrix_mul_tests.py (function extracted; the script's timing loop not run)
    src = open(os.path.join(REF, "matrix_mul_tests.py")).read()
    fn = [n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "ternary_mm"][0]
    ns = {"torch": torch}
    exec(compile(ast.Module([fn], []), "matrix_mul_tests.py", "exec"), ns)
    torch.manual_seed(5)
    a = torch.randn(6, 32)
    b = torch.randn(32, 32)
    b[torch.rand(32, 32) < 0.5] = 0.0
    g["tmm_a"], g["tmm_b"] = a.numpy(), b.numpy()
    g["tmm_out"] = ns["ternary_mm"](a, b).numpy()

    np.savez_compressed(os.path.join(OUT, "reference_outputs.npz"), **g)
    print("wrote", os.path.join(OUT, "reference_outputs.npz"), len(g), "arrays",
          os.path.getsize(os.path.join(OUT, "reference_outputs.npz")), "bytes")


if __name__ == "__main__":
    main()
