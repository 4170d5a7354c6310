"""A NumPy-backed graph-mode stand-in for the dozen ``tensorflow`` (1.x) ops the reference's model code calls.

TEST INFRASTRUCTURE ONLY (used by tests/golden/make_golden.py in the build container).  TensorFlow 1.14 is absent from
this image, so the reference's TF1 graph code cannot run as is; this module lets the reference's OWN files
(mopo/models/constructor.py, bnn.py, fc.py, utils.py -- imported unmodified from /root/reference) build their graph and
``sess.run`` it: every op records (function, inputs) and is evaluated lazily in float32 NumPy.  What this pins is the
reference's wiring (layer order, head doubling at bnn.py:185, which activation feeds the variance head at bnn.py:657-667,
the clamp formula at :654-655 / :669-670, the einsum / matmul split at fc.py:98-101, the non-factored aggregation at
bnn.py:261-263); the numerics of TensorFlow's own kernels (GEMM summation order, its softplus / sigmoid) stay third-party.
"""
import contextlib
import sys
import types

import numpy as np

float32 = np.float32


class _Dim(int):
    @property
    def value(self):
        return int(self)


class _Shape(tuple):
    @property
    def ndims(self):
        return len(self)

    def as_list(self):
        return [int(d) for d in self]


class Tensor:
    """Deferred op: fn(*evaluated inputs).  ``_dummy`` is the op evaluated with unknown (None) dims = 1, for static shapes."""

    def __init__(self, fn, inputs, name=None):
        self.fn, self.inputs, self.name = fn, list(inputs), name
        self._dummy = np.asarray(fn(*[_dummy(i) for i in self.inputs]))

    @property
    def shape(self):
        return _Shape(_Dim(d) for d in self._dummy.shape)

    def get_shape(self):
        return self.shape

    def _binary(self, other, f, rev=False):
        return Tensor((lambda a, b: f(b, a)) if rev else f, [self, other])

    def __add__(self, o): return self._binary(o, np.add)
    def __radd__(self, o): return self._binary(o, np.add, True)
    def __sub__(self, o): return self._binary(o, np.subtract)
    def __rsub__(self, o): return self._binary(o, np.subtract, True)
    def __mul__(self, o): return self._binary(o, np.multiply)
    def __rmul__(self, o): return self._binary(o, np.multiply, True)
    def __truediv__(self, o): return self._binary(o, np.divide)
    def __rtruediv__(self, o): return self._binary(o, np.divide, True)
    def __neg__(self): return Tensor(np.negative, [self])
    def __getitem__(self, idx): return Tensor(lambda a: a[idx], [self])
    __array_priority__ = 100


class Variable(Tensor):
    def __init__(self, initial_value, dtype=None, name=None, trainable=True):
        self.value = np.array(initial_value, dtype=np.float32 if dtype is None else dtype)
        self.fn, self.inputs, self.name, self.trainable = None, [], name, trainable

    @property
    def _dummy(self):
        return self.value

    def load(self, v, session=None):
        v = np.asarray(v, dtype=self.value.dtype)
        assert v.shape == self.value.shape, (v.shape, self.value.shape)
        self.value = v.copy()

    def eval(self, session=None):
        return self.value.copy()

    def assign(self, v):
        return Tensor(lambda x: self.load(x) or self.value, [v])


class Placeholder(Tensor):
    def __init__(self, dtype, shape, name=None):
        self.fn, self.inputs, self.name, self.dtype = None, [], name, dtype
        self._static = [1 if d is None else int(d) for d in shape]

    @property
    def _dummy(self):
        return np.zeros(self._static, dtype=self.dtype)


def _dummy(x):
    return x._dummy if isinstance(x, Tensor) else x


def _f32(x):
    """Python / NumPy constants enter the graph as float32, as TF converts them to the other operand's dtype."""
    a = np.asarray(x)
    return a.astype(np.float32) if a.dtype == np.float64 else a


def _evaluate(node, feed, memo):
    if not isinstance(node, Tensor):
        return _f32(node)
    if id(node) in memo:
        return memo[id(node)]
    if isinstance(node, Variable):
        out = node.value
    elif isinstance(node, Placeholder):
        out = np.asarray(feed[node], dtype=node.dtype)        # TF casts the fed array to the placeholder's dtype
    else:
        out = node.fn(*[_evaluate(i, feed, memo) for i in node.inputs])
    memo[id(node)] = out
    return out


class Session:
    def __init__(self, config=None, graph=None):
        pass

    def run(self, fetches, feed_dict=None):
        feed, memo = feed_dict or {}, {}
        if isinstance(fetches, (list, tuple)):
            return [None if f is None else np.array(_evaluate(f, feed, memo)) for f in fetches]
        return None if fetches is None else np.array(_evaluate(fetches, feed, memo))

    @contextlib.contextmanager
    def as_default(self):
        yield self


@contextlib.contextmanager
def variable_scope(name, *a, **k):
    yield


def placeholder(dtype, shape=None, name=None):
    return Placeholder(dtype, shape, name)


def get_variable(name, shape=None, initializer=None, trainable=True, dtype=None):
    return Variable(initializer(shape), name=name, trainable=trainable)


def constant_initializer(v):
    return lambda shape: np.full(shape, v, np.float32)


def truncated_normal_initializer(stddev=1.0, seed=0):
    def init(shape):
        rng = np.random.RandomState(abs(hash(tuple(shape))) % (2 ** 31))
        out = rng.standard_normal(size=shape)
        bad = np.abs(out) > 2.0
        while bad.any():
            out[bad] = rng.standard_normal(size=int(bad.sum()))
            bad = np.abs(out) > 2.0
        return (out * stddev).astype(np.float32)
    return init


def random_normal_initializer(*a, **k):
    return lambda shape: np.random.RandomState(0).standard_normal(size=shape).astype(np.float32)


def _op(f):
    return lambda *xs, **kw: Tensor(lambda *v: f(*v, **kw), xs)


def _sigmoid(x):
    with np.errstate(over='ignore'):
        return (np.float32(1.0) / (np.float32(1.0) + np.exp(-x))).astype(x.dtype)


def _softplus(x):
    return np.logaddexp(np.zeros((), x.dtype), x)


def _softmax(x):
    e = np.exp(x - x.max(-1, keepdims=True))
    return e / e.sum(-1, keepdims=True)


identity = _op(lambda x: x)
tanh = _op(np.tanh)
sigmoid = _op(_sigmoid)
exp = _op(np.exp)
square = _op(np.square)
multiply = lambda a, b, name=None: Tensor(np.multiply, [a, b])
matmul = _op(np.matmul)
transpose = _op(np.transpose)
reshape = lambda x, shape: Tensor(lambda a: a.reshape(shape), [x])
stop_gradient = identity
add_n = lambda xs: Tensor(lambda *v: sum(v[1:], v[0]), xs)


def einsum(eq, *xs):
    if eq == "ij,ajk->aik":          # fc.py:99: the shared-input ensemble contraction == a broadcast (BLAS) matmul
        return Tensor(lambda x, w: np.matmul(x[None], w), xs)
    return Tensor(lambda *v: np.einsum(eq, *v), xs)


def reduce_mean(x, axis=None, **k):
    return Tensor(lambda a: np.mean(a, axis=axis, dtype=a.dtype) if a.size else np.zeros(np.mean(np.zeros([max(d, 1) for d in a.shape]), axis=axis).shape, a.dtype), [x])


def reduce_sum(x, axis=None, **k):
    return Tensor(lambda a: np.sum(a, axis=axis, dtype=a.dtype), [x])


def variables_initializer(vs, name=None):
    return None


class _Optimizer:
    def __init__(self, **kw):
        self.kw = kw

    def minimize(self, loss, var_list=None):
        return None              # training is not exercised through the stub (oracle/train.py is pinned separately)

    def variables(self):
        return []


class GPUOptions:
    def __init__(self, **k):
        pass


class ConfigProto:
    def __init__(self, **k):
        pass


nn = types.SimpleNamespace(relu=_op(lambda x: np.maximum(x, 0)), softmax=_op(_softmax), softplus=_op(_softplus),
                           l2_loss=_op(lambda w: np.sum(np.square(w)) / 2), l2_normalize=_op(lambda x: x / np.linalg.norm(x)))
train = types.SimpleNamespace(AdamOptimizer=_Optimizer)


def install():
    """Register this module as ``tensorflow`` (and empty RLA stubs) so the reference's files import unmodified."""
    me = sys.modules[__name__]
    sys.modules['tensorflow'] = me
    rla = types.ModuleType('RLA')
    easy = types.ModuleType('RLA.easy_log')
    easy.logger = types.SimpleNamespace(ma_record_tabular=lambda *a, **k: None, dump_tabular=lambda *a, **k: None,
                                        record_tabular=lambda *a, **k: None, info=lambda *a, **k: None)
    tst = types.ModuleType('RLA.easy_log.tester')
    tst.tester = types.SimpleNamespace(time_step_holder=types.SimpleNamespace(set_time=lambda *a: None))
    rla.easy_log = easy
    easy.tester = tst
    sys.modules.update({'RLA': rla, 'RLA.easy_log': easy, 'RLA.easy_log.tester': tst})
    return me
