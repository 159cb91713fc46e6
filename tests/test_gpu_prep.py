"""GPU parity, weight preparation: CUDA (through the C ABI) vs the CPU oracle, bit-exact."""
import numpy as np
import pytest
import torch

from oracle import ref_int8, ref_modules as om, ref_packing as rp
from ternary_vllm_b200 import ops

pytestmark = pytest.mark.gpu


def latent(N, K, seed, dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(N, K, generator=g).clip(-1, 1)
    if K >= 6:
        W[0, :6] = torch.tensor([0.5, -0.5, 0.49999, -0.49999, 0.0, -0.0])
    return W.to(dtype)


@pytest.mark.parametrize("N,K", [(1, 1), (3, 5), (48, 100), (37, 259), (64, 64), (128, 4096), (300, 1027)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("mode", ["ternary", "binary", "binary01"])
def test_quantize_pack_bit_exact(cuda, N, K, dtype, mode):
    W = latent(N, K, N * 1000 + K, dtype)
    packed, codes = ops.quantize_pack(W.to(cuda), mode, return_codes=True)
    want = {"ternary": om.ternary_codes, "binary": om.sign_codes, "binary01": om.step_codes}[mode](W).numpy()
    assert np.array_equal(codes.cpu().numpy(), want)
    kb = (K + 3) // 4
    ref_bytes = rp.pack_C(want)                                   # reference format C
    got = packed.cpu().numpy()
    assert got.shape == (N, ops.packed_pitch(K))
    assert np.array_equal(got[:, :kb], ref_bytes)
    assert np.all(got[:, kb:] == 0x55)                            # padding = code 1 (weight 0)
    assert np.array_equal(ops.unpack(packed, K).cpu().numpy(), want)


def test_quantize_pack_strided_rows_and_tight_pitch(cuda):
    W = latent(20, 96, 5).to(cuda)
    view = W[:, :60]                                              # row pitch 96, K = 60
    packed = ops.quantize_pack(view, "ternary", ld_packed=15)     # tight pitch = the reference's shape
    want = rp.pack_C(om.ternary_codes(view.cpu()).numpy())
    assert np.array_equal(packed.cpu().numpy(), want)


def test_golden_weights_pack(cuda, golden):
    for tag in ("small", "k4", "ragged"):
        W = torch.from_numpy(golden[f"{tag}_W"])
        _, codes = ops.quantize_pack(W.to(cuda), "ternary", return_codes=True)
        assert np.array_equal(codes.cpu().numpy(), np.rint(golden[f"{tag}_q_ternary"]).astype(np.int8))
        _, codes = ops.quantize_pack(W.to(cuda), "binary", return_codes=True)
        assert np.array_equal(codes.cpu().numpy(), np.rint(golden[f"{tag}_q_sign"]).astype(np.int8))
        _, codes = ops.quantize_pack(W.to(cuda), "binary01", return_codes=True)
        assert np.array_equal(codes.cpu().numpy(), np.rint(golden[f"{tag}_q_step"]).astype(np.int8))


def test_pack_formats_known_answers_and_round_trip(cuda):
    t = torch.from_numpy(rp.KAT_T.astype(np.int8)).to(cuda)
    assert ops.pack_format(t, "A").cpu().numpy().tobytes() == rp.KAT_A
    assert ops.pack_format(t, "B").cpu().numpy().tobytes() == rp.KAT_B
    assert ops.pack_format(t.reshape(1, -1), "C").cpu().numpy().tobytes() == rp.KAT_C
    rng = np.random.default_rng(0)
    for T in (1, 5, 5031, 4096, 100003):
        v = rng.integers(-1, 2, size=T).astype(np.int8)
        for fmt, pk, un in (("A", rp.pack_A, rp.unpack_A), ("B", rp.pack_B, rp.unpack_B)):
            got = ops.pack_format(torch.from_numpy(v).to(cuda), fmt)
            assert np.array_equal(got.cpu().numpy(), pk(v)), (fmt, T)
            back = ops.unpack_format(got, fmt).cpu().numpy()
            assert np.array_equal(back, un(pk(v), np.int8)) and np.array_equal(back[:T], v)
    w = rng.integers(-1, 2, size=(33, 259)).astype(np.int8)
    got = ops.pack_format(torch.from_numpy(w).to(cuda), "C")
    assert np.array_equal(got.cpu().numpy(), rp.pack_C(w))
    assert np.array_equal(ops.unpack_format(got, "C", 33, 259).cpu().numpy(), w)
    # reference quirks: 0b11 decodes to 0 in A and +2 in B
    b = torch.tensor([0b11], dtype=torch.uint8, device=cuda)
    assert ops.unpack_format(b, "A").cpu()[0] == 0 and ops.unpack_format(b, "B").cpu()[0] == 2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_clip_latent(cuda, golden, dtype):
    for k in ("0", "1"):
        w = torch.from_numpy(golden["clip_in" + k]).to(dtype)
        got = ops.clip_latent_(w.clone().to(cuda)).cpu()
        assert torch.equal(got, om.clip_latent(w))
    if dtype == torch.float32:
        assert np.array_equal(ops.clip_latent_(torch.from_numpy(golden["clip_in0"]).to(cuda)).cpu().numpy(), golden["clip_out0"])
    big = (torch.randn(1 << 20) * 2).to(dtype)
    assert torch.equal(ops.clip_latent_(big.clone().to(cuda)).cpu(), big.clip(-1, 1))


CASES = [("ternary", "none", "uniform"), ("ternary", "none", "none"), ("ternary", "none", "independent"),
         ("binary", "none", "uniform"), ("binary", "none", "independent"), ("binary", "std", "independent"),
         ("binary", "natural", "independent"), ("binary", 0.5, "uniform"),
         ("binary01", "none", "uniform"), ("binary01", "none", "independent"), ("binary01", "std", "uniform")]


def _gpu_init(W, qn, sc, ns):
    if qn == "ternary":
        return ops.ternary_init(W, per_row=ns != "uniform")
    pre, natural = 1.0, False
    if isinstance(sc, float):
        pre = sc
    elif sc == "std":
        pre = float(1.0 / W.std())
    elif sc == "natural":
        natural = True
    return ops.binary_init(W, per_row=ns == "independent", positive_only=qn == "binary01", prescale=pre, natural=natural)


@pytest.mark.parametrize("qn,sc,ns", CASES)
def test_scale_init_golden(cuda, golden, qn, sc, ns):
    W = torch.from_numpy(golden["l2q_W"]).to(cuda)
    latent_, scale = _gpu_init(W, qn, sc, ns)
    key = f"l2q_{qn}_{sc}_{ns}"
    # reductions run in fp64 in a fixed order on the GPU and in fp32 in torch: agree to ~1 ulp
    assert np.allclose(scale.cpu().numpy(), golden[key + "_scale"], rtol=2e-6, atol=0)
    assert np.allclose(latent_.cpu().numpy(), golden[key + "_latent"], rtol=2e-6, atol=1e-7)


@pytest.mark.parametrize("N,K", [(64, 2048), (11, 333)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_ternary_init_vs_oracle(cuda, N, K, dtype):
    g = torch.Generator().manual_seed(N + K)
    W = (torch.randn(N, K, generator=g) * 0.05).to(dtype)
    for ns in ("uniform", "none"):
        lat_o, sc_o = om.ternary_init(W, ns)
        lat, sc = ops.ternary_init(W.to(cuda), per_row=ns != "uniform")
        assert np.allclose(sc.cpu().numpy(), sc_o.float().numpy(), rtol=1e-5)
        assert np.allclose(lat.cpu().numpy(), lat_o.float().numpy(), rtol=1e-5, atol=1e-6)
        # codes agree except (rarely) for latent values within rounding of the +-0.5 threshold
        c_o = om.ternary_codes(lat_o.float())
        c = om.ternary_codes(lat.cpu())
        near = (lat_o.float().abs() - 0.5).abs() < 1e-5
        assert torch.equal(c[~near], c_o[~near])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("layout", [0, 1, 2, 3])
@pytest.mark.parametrize("M,K", [(1, 64), (5, 77), (3, 4096), (40, 1000), (2, 11008),
                                 # >= 4 rows per SM: the warp-per-row kernels (specialised for K = 1024 / 2048 / 4096, generic otherwise)
                                 (700, 4096), (650, 1024), (600, 2048), (640, 512), (610, 3000)])
def test_quant_act_bit_exact(cuda, dtype, layout, M, K):
    g = torch.Generator().manual_seed(M * 7 + K)
    x = (torch.randn(M, K, generator=g) * 3).to(dtype)
    if M > 2:
        x[1] = 0
    q, s, sq = ops.quant_act_i8(x.to(cuda), layout=layout)
    qo, so, sqo = ref_int8.quant_act(x.float().numpy())
    assert np.array_equal(s.cpu().numpy(), so) and np.array_equal(sq.cpu().numpy(), sqo)
    ldq = q.shape[1]
    qpad = np.zeros((M, ldq), np.int8)
    qpad[:, :K] = qo
    if layout == 1:
        qpad = qpad[:, ref_int8.kperm_index(ldq)]
    elif layout == 2:
        assert ldq == (K + 4095) // 4096 * 4096
        qpad = qpad[:, ref_int8.gemv_layout_index(K)]
    elif layout == 3:
        assert ldq == (K + 255) // 256 * 256
        qpad = qpad[:, ref_int8.split_layout_index(ldq)]
    assert np.array_equal(q.cpu().numpy(), qpad)


# ---- the reference's own pack / unpack / packed-layer outputs (oracle/make_golden_packing.py) -----------------
@pytest.fixture(scope="module")
def pk():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    return np.load(os.path.join(root, "tests", "golden", "packing_vectors.npz"), allow_pickle=False)


@pytest.mark.parametrize("T", [1, 2, 3, 4, 5, 8, 63, 64, 1000, 5031])
def test_pack_formats_A_B_vs_reference_bytes(cuda, pk, T):
    t = torch.from_numpy(pk[f"ab_t_{T}"].astype(np.int8)).to(cuda)
    for fmt, key in (("A", "a"), ("B", "b")):
        got = ops.pack_format(t, fmt)
        assert np.array_equal(got.cpu().numpy(), pk[f"{key}_packed_{T}"]), (fmt, T)
        back = ops.unpack_format(got, fmt).cpu().numpy()
        assert np.array_equal(back, pk[f"{key}_unpacked_{T}"].astype(np.int8)), (fmt, T)


def test_unpack_formats_every_byte_vs_reference(cuda, pk):
    allb = torch.arange(256, dtype=torch.int32).to(torch.uint8).to(cuda)
    assert np.array_equal(ops.unpack_format(allb, "A").cpu().numpy(), pk["a_unpacked_allbytes"].astype(np.int8))
    assert np.array_equal(ops.unpack_format(allb, "B").cpu().numpy(), pk["b_unpacked_allbytes"].astype(np.int8))


@pytest.mark.parametrize("tag", ["c_5x4", "c_7x8", "c_9x32", "c_16x64", "c_12x260"])
def test_format_C_ops_and_fused_op_vs_reference_packed_layers(cuda, pk, tag):
    """a14/a16/a17: pack_fn bytes, unpack, and ops.ternary_matmul's maths against the outputs of the reference's own
    ScaledTernaryPackedMM.pack_fn / .call and ScaledTernaryTf.call (float activations: tb200_gemv_f32)."""
    w, x = pk[tag + "_w"], pk[tag + "_x"]
    N, K = w.shape
    wt = torch.from_numpy(w).to(cuda)
    assert np.array_equal(ops.pack_format(wt, "C").cpu().numpy(), pk[tag + "_packed"])               # ops.pack_fn(int8 [N,K])
    internal = ops.quantize_pack(wt.float(), "ternary")                                              # ternarise + pack
    assert np.array_equal(internal.cpu().numpy()[:, : (K + 3) // 4], pk[tag + "_packed"])
    assert np.array_equal(ops.unpack(internal, K).cpu().numpy(), w)                                  # ops.unpack_ternary
    scale, bias = torch.from_numpy(pk[tag + "_scale"]).to(cuda), torch.from_numpy(pk[tag + "_bias"]).to(cuda)
    for clip in (100.0, 0.5):
        y = ops.gemv_f32(torch.from_numpy(x).to(cuda), internal, scale, bias, clip).cpu().numpy()
        want = pk[tag + f"_mm_clip{clip}"]
        assert np.allclose(y, want, rtol=1e-5, atol=2e-6), np.abs(y - want).max()
    y = ops.gemv_f32(torch.from_numpy(x).to(cuda), internal, scale, bias, 100.0).cpu().numpy()
    assert np.allclose(y, pk[tag + "_tf_call"], rtol=1e-5, atol=2e-6)
    x3 = pk[tag + "_x3"]
    y3 = ops.gemv_f32(torch.from_numpy(x3.reshape(-1, K)).to(cuda), internal, scale, bias, 100.0).cpu().numpy()
    assert np.allclose(y3.reshape(2, 3, N), pk[tag + "_mm3"], rtol=1e-5, atol=2e-6)


def test_ternary_mm_golden_on_gpu(cuda, golden):
    """a18: matrix_mul_tests.py:29-37 `ternary_mm(m, tm)` == m @ sign(tm)^T, through tb200_gemv_f32 (scale 1, no bias,
    no effective clip) on weights whose codes are sign(tm) in {-1,0,+1}."""
    a, b = golden["tmm_a"], golden["tmm_b"]
    codes = np.sign(b).astype(np.float32)                       # tm[i] > 0 -> +1, < 0 -> -1, == 0 -> skipped
    packed = ops.quantize_pack(torch.from_numpy(codes).to(cuda), "ternary")
    y = ops.gemv_f32(torch.from_numpy(a).to(cuda), packed, None, None, 3.0e38).cpu().numpy()
    assert np.allclose(y, golden["tmm_out"], rtol=1e-5, atol=1e-5)
