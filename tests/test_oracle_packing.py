"""Pins oracle/ref_packing.py: golden vectors produced by EXECUTING the reference's own pack / unpack source
(oracle/make_golden_packing.py: impl/tf/moondream/ternary.py run with oracle/tf_stub.py standing in for TensorFlow),
known-answer bytes and the round trip the reference itself tests (ternary_tests.py:17-38)."""
import os

import numpy as np
import pytest

from oracle import ref_packing as rp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
AB_LENGTHS = (1, 2, 3, 4, 5, 8, 63, 64, 1000, 5031)
C_TAGS = ("c_5x4", "c_7x8", "c_9x32", "c_16x64", "c_12x260")


@pytest.fixture(scope="module")
def pk():
    return np.load(os.path.join(ROOT, "tests", "golden", "packing_vectors.npz"), allow_pickle=False)


@pytest.mark.parametrize("T", AB_LENGTHS)
def test_formats_A_B_match_reference_source(pk, T):
    t = pk[f"ab_t_{T}"]
    assert np.array_equal(rp.pack_A(t), pk[f"a_packed_{T}"])
    assert np.array_equal(rp.pack_B(t), pk[f"b_packed_{T}"])
    assert np.array_equal(rp.unpack_A(pk[f"a_packed_{T}"]), pk[f"a_unpacked_{T}"])
    assert np.array_equal(rp.unpack_B(pk[f"b_packed_{T}"]), pk[f"b_unpacked_{T}"])
    assert np.array_equal(pk[f"b_unpacked_ec_{T}"], pk[f"b_unpacked_{T}"])      # ternary.py:159-175 == :136-156
    assert np.array_equal(pk[f"a_unpacked_{T}"][:T], t) and np.array_equal(pk[f"b_unpacked_{T}"][:T], t)


def test_reference_decodes_every_byte(pk):
    allb = np.arange(256, dtype=np.uint8)
    assert np.array_equal(rp.unpack_A(allb), pk["a_unpacked_allbytes"])
    assert np.array_equal(rp.unpack_B(allb), pk["b_unpacked_allbytes"])
    assert np.array_equal(rp.pack_B(pk["b_w2d"]), pk["b_packed_2d"])            # 2-D input is flattened (ternary.py:115)


@pytest.mark.parametrize("tag", C_TAGS)
def test_format_C_and_packed_mm_match_reference_source(pk, tag):
    w, x, scale, bias = pk[tag + "_w"], pk[tag + "_x"], pk[tag + "_scale"], pk[tag + "_bias"]
    assert np.array_equal(rp.pack_C(w), pk[tag + "_packed"])
    assert np.array_equal(rp.unpack_C(pk[tag + "_packed"], w.shape[1]), w)
    for clip in (100.0, 0.5):
        want = pk[tag + f"_mm_clip{clip}"]
        got = np.clip(rp.packed_mm_C(x, pk[tag + "_packed"]) * scale + bias, -clip, clip)
        assert np.allclose(got, want, rtol=0, atol=1e-6)
    # the reference's two formulations of the packed layer (bit-plane matmuls vs unpack -> matmul) agree
    assert np.allclose(pk[tag + "_tf_call"], pk[tag + "_mm_clip100.0"], rtol=0, atol=2e-6)
    assert pk[tag + "_mm3"].shape == (2, 3, w.shape[0])


def test_reference_kat_and_ragged_quirk(pk):
    assert pk["kat_a"].tobytes() == rp.KAT_A and pk["kat_b"].tobytes() == rp.KAT_B and pk["kat_c"].tobytes() == rp.KAT_C
    # for K % 4 != 0 the reference's 2-D pack_fn raises (rank mismatch in tf.concat, ternary.py:380); padding such rows
    # with code 1 (value 0) is this repo's documented extension
    assert int(pk["c_ragged_pack_raises"]) == 1


def test_live_reference_source_reproduces_golden(pk):
    """Build container only: re-execute the reference source and compare with the committed vectors."""
    if not os.path.isdir("/root/reference/impl"):
        pytest.skip("/root/reference not present (GPU box)")
    from oracle import make_golden_packing as mg
    live = mg.generate()
    assert sorted(live) == sorted(pk.files)
    for k in pk.files:
        assert np.array_equal(np.asarray(live[k]), pk[k]), k


def test_known_answer_bytes():
    assert rp.pack_A(rp.KAT_T).tobytes() == rp.KAT_A
    assert rp.pack_B(rp.KAT_T).tobytes() == rp.KAT_B
    assert rp.pack_C(rp.KAT_T.reshape(1, -1)).tobytes() == rp.KAT_C


@pytest.mark.parametrize("fmt", ["A", "B"])
def test_reference_round_trip_5031(fmt):
    # ternary_tests.py draws tf.random.uniform((5031,), -1, 1, int32), i.e. values in {-1, 0}
    rng = np.random.default_rng(0)
    t = rng.integers(-1, 1, size=5031).astype(np.int32)
    pack, unpack = (rp.pack_A, rp.unpack_A) if fmt == "A" else (rp.pack_B, rp.unpack_B)
    packed = pack(t)
    assert packed.shape == ((5031 + 3) // 4,)
    back = unpack(packed)[: t.shape[0]]
    assert np.all(back == t)


@pytest.mark.parametrize("fmt", ["A", "B"])
def test_round_trip_full_alphabet(fmt):
    rng = np.random.default_rng(1)
    t = rng.integers(-1, 2, size=4099).astype(np.int32)
    pack, unpack = (rp.pack_A, rp.unpack_A) if fmt == "A" else (rp.pack_B, rp.unpack_B)
    assert np.all(unpack(pack(t))[: t.shape[0]] == t)


@pytest.mark.parametrize("K", [1, 3, 4, 5, 64, 259])
def test_format_C_round_trip_and_padding(K):
    rng = np.random.default_rng(K)
    w = rng.integers(-1, 2, size=(7, K)).astype(np.int8)
    p = rp.pack_C(w)
    assert p.shape == (7, (K + 3) // 4)
    assert np.all(rp.unpack_C(p, K) == w)
    if K % 4:
        # padding fields hold code 1 (value 0): ternary.py:380-386 pads with zeros before +1
        last = p[:, -1]
        for i in range(K % 4, 4):
            assert np.all(((last >> (2 * i)) & 3) == 1)


def test_packed_mm_matches_dense():
    rng = np.random.default_rng(3)
    w = rng.integers(-1, 2, size=(9, 30)).astype(np.int8)
    x = rng.standard_normal((4, 30)).astype(np.float32)
    got = rp.packed_mm_C(x, rp.pack_C(w))
    assert np.allclose(got, x @ w.T.astype(np.float32), atol=1e-5)


def test_format_quirks():
    # code 0b11: format A decodes to 0, format B to +2 (the reference guards neither)
    assert rp.unpack_A(np.array([0b11], np.uint8))[0] == 0
    assert rp.unpack_B(np.array([0b11], np.uint8))[0] == 2
