"""Pins oracle/ref_packing.py: known-answer bytes and the round trip the
reference itself tests (impl/tf/moondream/ternary_tests.py:17-38)."""
import numpy as np
import pytest

from oracle import ref_packing as rp


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
