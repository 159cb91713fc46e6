"""Generate tests/golden/packing_vectors.npz by EXECUTING THE REFERENCE'S OWN pack / unpack source
(/root/reference/impl/tf/moondream/ternary.py) with ``oracle/tf_stub.py`` bound to the name ``tf``.

    python oracle/make_golden_packing.py          # rewrites tests/golden/packing_vectors.npz

TensorFlow, larq and the author's Larq-Compute-Engine fork are not installed (and cannot be), so
``ternary.py`` cannot be imported.  Its functions are therefore taken from the file with ``ast`` (the
source text is compiled where it lies; nothing is copied into this repo) and run on seeded inputs:

  pack_ternary / unpack_ternary_iterative                      (format A, ternary.py:57-78,101-110)
  pack_ternary2 / unpack_ternary_iterative2(_early_concat)     (format B, :114-175)
  ScaledTernaryPackedMM.pack_fn / .call                        (format C + bit-plane matmul, :374-398,485-538)
  ScaledTernaryTf.call                                         (unpack -> matmul -> scale/bias/clip, :210-250)
  TernaryCase.test_packing_unpacking                           (the reference's own test, ternary_tests.py:17-27)

``tests/test_oracle_packing.py`` re-runs this extraction live when /root/reference is present and
compares with the committed vectors; on the GPU box only the vectors travel.
"""
import ast
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import tf_stub as tf  # noqa: E402

REF = "/root/reference"
TERNARY_PY = os.path.join(REF, "impl", "tf", "moondream", "ternary.py")
TESTS_PY = os.path.join(REF, "impl", "tf", "moondream", "ternary_tests.py")
OUT = os.path.join(HERE, "..", "tests", "golden", "packing_vectors.npz")

FUNCS = ["pack_ternary", "unpack_ternary_iterative", "pack_ternary2", "unpack_ternary_iterative2",
         "unpack_ternary_iterative2_early_concat"]
METHODS = {"ScaledTernaryPackedMM": ["pack_fn", "call"], "ScaledTernaryTf": ["call"]}


def _compile(nodes, filename, ns):
    for n in nodes:
        n.decorator_list = []                 # @tf.function: eager execution is the same computation
    exec(compile(ast.Module(body=list(nodes), type_ignores=[]), filename, "exec"), ns)


def load_reference_functions():
    """name -> callable compiled from the reference source (functions and selected methods)."""
    tree = ast.parse(open(TERNARY_PY).read())
    ns = {"tf": tf, "np": np}
    top = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in FUNCS]
    assert sorted(n.name for n in top) == sorted(FUNCS)
    _compile(top, TERNARY_PY, ns)
    out = {name: ns[name] for name in FUNCS}
    for cls, names in METHODS.items():
        cdef = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls][0]
        for name in names:
            fn = [n for n in cdef.body if isinstance(n, ast.FunctionDef) and n.name == name][0]
            sub = {"tf": tf, "np": np}
            _compile([fn], TERNARY_PY, sub)
            out[f"{cls}.{name}"] = sub[name]
    return out


def run_reference_test_case(fns):
    """The reference's own TernaryCase.test_packing_unpacking (ternary_tests.py:17-27), for A and B."""
    import unittest
    tree = ast.parse(open(TESTS_PY).read())
    cdef = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "TernaryCase"][0]
    ns = {"tf": tf, "TestCase": unittest.TestCase, "pack_ternary": fns["pack_ternary"],
          "unpack_ternary_iterative": fns["unpack_ternary_iterative"]}
    _compile([cdef], TESTS_PY, ns)
    for pk, up in (("pack_ternary", "unpack_ternary_iterative"), ("pack_ternary2", "unpack_ternary_iterative2")):
        case = ns["TernaryCase"](pack_fn=fns[pk], unpack_fn=fns[up], methodName="test_packing_unpacking")
        case.test_packing_unpacking()          # raises AssertionError on failure
    return True


def generate():
    fns = load_reference_functions()
    assert run_reference_test_case(fns)
    g = {}
    rng = np.random.default_rng(20261104)
    # ---- formats A and B: 1-D, ragged lengths, the full alphabet {-1,0,+1} ------------------------------
    for T in (1, 2, 3, 4, 5, 8, 63, 64, 1000, 5031):
        t = rng.integers(-1, 2, size=T).astype(np.int32)
        g[f"ab_t_{T}"] = t
        pa = fns["pack_ternary"](tf.constant(t)).numpy()
        pb = fns["pack_ternary2"](tf.constant(t)).numpy()
        g[f"a_packed_{T}"], g[f"b_packed_{T}"] = pa, pb
        g[f"a_unpacked_{T}"] = fns["unpack_ternary_iterative"](tf.constant(pa)).numpy()
        g[f"b_unpacked_{T}"] = fns["unpack_ternary_iterative2"](tf.constant(pb)).numpy()
        g[f"b_unpacked_ec_{T}"] = fns["unpack_ternary_iterative2_early_concat"](tf.constant(pb)).numpy()
    # int8 input, as conversion.py:52-53 feeds it (cast to int8, flattened [N*K])
    w2 = rng.integers(-1, 2, size=(6, 10)).astype(np.int8)
    g["b_w2d"] = w2
    g["b_packed_2d"] = fns["pack_ternary2"](tf.constant(w2)).numpy()
    # every byte value through the unpackers (0b11 fields: A -> 0, B -> +2)
    allb = np.arange(256, dtype=np.uint8)
    g["a_unpacked_allbytes"] = fns["unpack_ternary_iterative"](tf.constant(allb)).numpy()
    g["b_unpacked_allbytes"] = fns["unpack_ternary_iterative2"](tf.constant(allb)).numpy()
    # known-answer vector of SURVEY.md §8c
    kat = np.array([1, -1, 0, 1, 0, 0, -1, 1], dtype=np.int32)
    g["kat_t"] = kat
    g["kat_a"] = fns["pack_ternary"](tf.constant(kat)).numpy()
    g["kat_b"] = fns["pack_ternary2"](tf.constant(kat)).numpy()
    g["kat_c"] = fns["ScaledTernaryPackedMM.pack_fn"](None, tf.constant(kat.reshape(1, -1).astype(np.int8))).numpy()

    # ---- format C (pack_fn) and the bit-plane matmul (call), ragged K -----------------------------------
    # K % 4 == 0 only: for a ragged K the reference's pack_fn raises (tf.concat of the rank-2 weights with a rank-1
    # zero pad, ternary.py:380) -- recorded below; every layer the reference converts has K in {2048, 8192}
    try:
        fns["ScaledTernaryPackedMM.pack_fn"](None, tf.constant(rng.integers(-1, 2, size=(3, 5)).astype(np.int8)))
        g["c_ragged_pack_raises"] = np.array(0)
    except ValueError:
        g["c_ragged_pack_raises"] = np.array(1)
    for (N, K, M) in ((5, 4, 2), (7, 8, 3), (9, 32, 4), (16, 64, 1), (12, 260, 5)):
        w = rng.integers(-1, 2, size=(N, K)).astype(np.int8)
        x = rng.standard_normal((M, K)).astype(np.float32)
        scale = (rng.random(N) * 0.2 + 0.01).astype(np.float32)
        bias = rng.standard_normal(N).astype(np.float32)
        tag = f"c_{N}x{K}"
        g[tag + "_w"], g[tag + "_x"], g[tag + "_scale"], g[tag + "_bias"] = w, x, scale, bias
        pw = fns["ScaledTernaryPackedMM.pack_fn"](None, tf.constant(w)).numpy()
        g[tag + "_packed"] = pw
        for clip in (100.0, 0.5):
            me = types.SimpleNamespace(pw=tf.constant(pw), orig_shape=(N, K), scale=tf.constant(scale),
                                       bias=tf.constant(bias), clip_val=tf.abs(clip), act=None)
            g[tag + f"_mm_clip{clip}"] = fns["ScaledTernaryPackedMM.call"](me, tf.constant(x)).numpy()
        # 3-D input (batch, seq, K), as the Keras model calls it
        x3 = rng.standard_normal((2, 3, K)).astype(np.float32)
        me = types.SimpleNamespace(pw=tf.constant(pw), orig_shape=(N, K), scale=tf.constant(scale),
                                   bias=tf.constant(bias), clip_val=tf.abs(100.0), act=None)
        g[tag + "_x3"] = x3
        g[tag + "_mm3"] = fns["ScaledTernaryPackedMM.call"](me, tf.constant(x3)).numpy()
        # ScaledTernaryTf.call: format-B packed flat weights -> unpack -> reshape -> matmul -> scale, bias, clip
        pb = fns["pack_ternary2"](tf.constant(w.reshape(-1))).numpy()
        me = types.SimpleNamespace(pw=tf.constant(pb), orig_shape=(N, K), tot_size=N * K, scale=tf.constant(scale),
                                   bias=tf.constant(bias), clip_val=tf.abs(100.0), act=None, debug=False,
                                   unpack_fn=fns["unpack_ternary_iterative2"])
        g[tag + "_tf_call"] = fns["ScaledTernaryTf.call"](me, tf.constant(x)).numpy()
    return g


def main():
    g = generate()
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, len(g), "arrays", os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
