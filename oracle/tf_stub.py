"""A numpy-backed stand-in for the handful of TensorFlow names the reference's pack / unpack code
uses.  TEST INFRASTRUCTURE ONLY (nothing under ternary_vllm_b200/ imports it).

Purpose: TensorFlow is not installed here, so the reference's own packing source
(``impl/tf/moondream/ternary.py:57-175,374-398,485-538``) could not be executed in round 1 and
``oracle/ref_packing.py`` was a restatement nobody had run the reference against.  With this module
bound to the name ``tf``, ``oracle/make_golden_packing.py`` executes THE REFERENCE'S SOURCE TEXT
(extracted with ``ast``, not copied) and records its outputs as golden vectors.

Only the semantics those functions rely on are provided, each matching TensorFlow 2.10 (the version
``requirements.txt:10`` pins):

* tensors are immutable arrays with a static ``.shape`` (a tuple whose comparison with an int ``n``
  means ``== (n,)``, like ``TensorShape``), ``.dtype``, numpy-style slicing, ``==``, ``+ - * @ %``;
* ``tf.cast`` truncates like a C cast (bool -> 0/1, int -> uint8 modulo 256);
* ``tf.bitwise.*`` operate in the dtype of the tensor operand; a Python-int shift count is
  converted to that dtype first (what ``tf.convert_to_tensor`` does for the second argument);
* ``tf.concat`` / ``tf.stack`` / ``tf.reshape`` / ``tf.transpose`` / ``tf.pad`` / ``tf.zeros`` /
  ``tf.constant`` / ``tf.shape`` / ``tf.cond`` / ``tf.matmul(transpose_b=)`` /
  ``tf.clip_by_value`` / ``tf.abs`` / ``tf.print`` (no-op) / ``tf.random.uniform`` (integer
  dtypes: half-open ``[minval, maxval)`` like TF).
"""
from __future__ import annotations

import numpy as np

uint8 = np.dtype("uint8")
int8 = np.dtype("int8")
int32 = np.dtype("int32")
int64 = np.dtype("int64")
float32 = np.dtype("float32")
float16 = np.dtype("float16")


class _Shape(tuple):
    """TensorShape: a tuple of ints; ``shape == 1`` is ``shape == (1,)`` (TensorShape(1) is rank 1)."""

    def __eq__(self, other):
        if isinstance(other, (int, np.integer)):
            other = (int(other),)
        return tuple(self) == tuple(other)

    def __ne__(self, other):
        return not self.__eq__(other)

    __hash__ = tuple.__hash__

    def __getitem__(self, i):
        r = tuple.__getitem__(self, i)
        return _Shape(r) if isinstance(r, tuple) else r


def _raw(v):
    return v.a if isinstance(v, Tensor) else v


class Tensor:
    __array_priority__ = 100

    def __init__(self, a):
        self.a = np.asarray(a)

    shape = property(lambda self: _Shape(self.a.shape))
    dtype = property(lambda self: self.a.dtype)

    def numpy(self):
        return self.a

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            idx = tuple(_raw(i) for i in idx)
        else:
            idx = _raw(idx)
        return Tensor(self.a[idx])

    def _bin(self, other, fn, reflected=False):
        o = _raw(other)
        if not isinstance(o, np.ndarray):
            o = np.asarray(o, dtype=self.a.dtype)      # Python scalars take the tensor's dtype (tf.convert_to_tensor)
        return Tensor(fn(o, self.a) if reflected else fn(self.a, o))

    def __add__(self, o): return self._bin(o, np.add)
    def __radd__(self, o): return self._bin(o, np.add, True)
    def __sub__(self, o): return self._bin(o, np.subtract)
    def __rsub__(self, o): return self._bin(o, np.subtract, True)
    def __mul__(self, o): return self._bin(o, np.multiply)
    def __rmul__(self, o): return self._bin(o, np.multiply, True)
    def __mod__(self, o): return self._bin(o, np.mod)
    def __matmul__(self, o): return Tensor(np.matmul(self.a, _raw(o)))
    def __neg__(self): return Tensor(-self.a)
    def __eq__(self, o): return self._bin(o, np.equal)
    def __ne__(self, o): return self._bin(o, np.not_equal)
    def __lt__(self, o): return self._bin(o, np.less)
    def __bool__(self): return bool(self.a)
    def __int__(self): return int(self.a)
    def __index__(self): return int(self.a)
    __hash__ = None


def _t(v, dtype=None):
    return Tensor(np.asarray(_raw(v), dtype=dtype))


def zeros(shape, dtype=float32):
    return Tensor(np.zeros(tuple(int(s) for s in shape) if not isinstance(shape, (int, np.integer)) else (int(shape),), dtype=dtype))


def constant(value, dtype=None):
    return _t(value, dtype)


convert_to_tensor = constant


def cast(x, dtype):
    return Tensor(np.asarray(_raw(x)).astype(dtype))


def concat(values, axis):
    return Tensor(np.concatenate([np.asarray(_raw(v)) for v in values], axis=axis))


def stack(values, axis=0):
    return Tensor(np.stack([np.asarray(_raw(v)) for v in values], axis=axis))


def reshape(tensor, shape):
    return Tensor(np.reshape(_raw(tensor), tuple(int(s) for s in shape)))


def transpose(a):
    return Tensor(np.transpose(_raw(a)))


def shape(x):
    return Tensor(np.asarray(np.asarray(_raw(x)).shape, dtype=np.int32))


def pad(tensor, paddings):
    return Tensor(np.pad(_raw(tensor), [(int(_raw(a)), int(_raw(b))) for a, b in paddings]))


def cond(pred, true_fn, false_fn):
    return true_fn() if bool(_raw(pred)) else false_fn()


def matmul(a, b, transpose_b=False):
    b = np.asarray(_raw(b))
    return Tensor(np.matmul(_raw(a), np.swapaxes(b, -1, -2) if transpose_b else b))


def clip_by_value(t, lo, hi):
    return Tensor(np.clip(_raw(t), _raw(lo), _raw(hi)))


def abs(x):  # noqa: A001 - mirrors tf.abs
    return _t(np.abs(_raw(x)))


def reduce_all(x):
    return Tensor(np.all(_raw(x)))


def print(*args, **kwargs):  # noqa: A001 - tf.print; the reference only calls it behind `self.debug`
    return None


def function(fn=None, **kwargs):
    """@tf.function: eager execution is the same computation."""
    return fn if fn is not None else (lambda f: f)


class _Bitwise:
    @staticmethod
    def _shift_operand(x, y):
        x = np.asarray(_raw(x))
        y = _raw(y)
        if not isinstance(y, np.ndarray):
            y = np.asarray(y, dtype=x.dtype)
        return x, y

    @classmethod
    def left_shift(cls, x, y):
        x, y = cls._shift_operand(x, y)
        return Tensor(np.left_shift(x, y).astype(x.dtype))

    @classmethod
    def right_shift(cls, x, y):
        x, y = cls._shift_operand(x, y)
        return Tensor(np.right_shift(x, y).astype(x.dtype))

    @staticmethod
    def _two(x, y):
        x, y = _raw(x), _raw(y)
        if not isinstance(x, np.ndarray):
            x = np.asarray(x, dtype=np.asarray(y).dtype)
        if not isinstance(y, np.ndarray):
            y = np.asarray(y, dtype=x.dtype)
        return x, y

    @classmethod
    def bitwise_or(cls, x, y):
        x, y = cls._two(x, y)
        return Tensor(np.bitwise_or(x, y))

    @classmethod
    def bitwise_and(cls, x, y):
        x, y = cls._two(x, y)
        return Tensor(np.bitwise_and(x, y))

    @classmethod
    def bitwise_xor(cls, x, y):
        x, y = cls._two(x, y)
        return Tensor(np.bitwise_xor(x, y))


bitwise = _Bitwise


class _Random:
    def __init__(self):
        self.rng = np.random.default_rng(1234)

    def set_seed(self, seed):
        self.rng = np.random.default_rng(seed)

    def uniform(self, shape, minval=0, maxval=None, dtype=float32):
        dtype = np.dtype(dtype)
        if dtype.kind in "iu":
            return Tensor(self.rng.integers(minval, maxval, size=tuple(shape)).astype(dtype))   # [minval, maxval)
        return Tensor(self.rng.uniform(minval, 1.0 if maxval is None else maxval, size=tuple(shape)).astype(dtype))


random = _Random()
