"""A TensorFlow-1.x graph-mode stand-in, just large enough to EXECUTE the reference's own model files.

TEST INFRASTRUCTURE ONLY (used by `tests/golden/make_ref_golden.py`, in the build container, where
/root/reference exists).  TensorFlow 1.x cannot be installed here, so the reference's arithmetic cannot
be run as shipped.  What CAN be run is everything the reference itself wrote: the graph construction
in `models/*.py` (which variable is looked up with which index tensor, what is concatenated in which
order, which weights are regularised, which tensors the training loop fetches) and its `fit` /
`evaluate` loops.  `install()` registers a module named `tensorflow` whose symbols build a lazy graph of
closures over torch-CPU float64 tensors; `Session.run` evaluates it, `AdamOptimizer.minimize` takes
gradients with autograd and applies TF1's documented Adam (`m`, `v`, bias correction through the
`beta*_power` accumulators, epsilon outside the square root, every row of a looked-up variable decayed:
`_apply_sparse_shared` assigns `m*beta1` / `v*beta2` to the whole slot before the scatter-add).

Only the ~50 symbols `models/*.py` touch are defined; anything else raises AttributeError, so a model
file cannot silently take a path this file does not restate.  Values of random initialisers do NOT
match TF's generators (they cannot); fixtures record the initial variables instead.
"""
from __future__ import annotations

import math
import sys
import types

import numpy as np
import torch

# arithmetic type of the stand-in: float64 (the fixtures; order-free truth) or, with TF1_SHIM_FLOAT=float32, TF's own
# float32 -- used ONLY to measure how far float32 arithmetic itself drifts from the float64 fixtures (the "fp32 floor"
# recorded beside them by make_ref_golden.py)
import os as _os
F = torch.float32 if _os.environ.get("TF1_SHIM_FLOAT", "float64") == "float32" else torch.float64


class _DType:
    def __init__(self, name, tt, is_float):
        self.name, self.tt, self.is_float = name, tt, is_float

    def __repr__(self):
        return "tf." + self.name


float32 = _DType("float32", F, True)
float64 = _DType("float64", F, True)
int32 = _DType("int32", torch.int64, False)
int64 = _DType("int64", torch.int64, False)


def _as_dtype(d):
    if d is None or isinstance(d, _DType):
        return d
    if d is float or d in (np.float32, np.float64) or str(d) in ("float32", "float64"):
        return float32
    if d is int or d in (np.int32, np.int64) or str(d) in ("int32", "int64"):
        return int32
    raise TypeError("dtype %r not handled by the TF1 stand-in" % (d,))


class TensorShape:
    def __init__(self, dims):
        self.dims = None if dims is None else list(dims)

    def as_list(self):
        if self.dims is None:
            raise ValueError("as_list() is not defined on an unknown TensorShape")
        return list(self.dims)

    def __getitem__(self, i):
        return self.as_list()[i]

    def __len__(self):
        return len(self.as_list())

    def __repr__(self):
        return "TensorShape(%r)" % (self.dims,)


class _Graph:
    def __init__(self):
        self.variables = []          # creation order (global variables)
        self.seed = None
        self.n_random = 0
        self.by_name = {}
        self.scope = []
        self.optimizer_slots = []

    def unique(self, name):
        base = "/".join(self.scope + [name]) if self.scope else name
        n, k = base, 0
        while n in self.by_name:
            k += 1
            n = "%s_%d" % (base, k)
        return n


_graph = _Graph()


def Graph():
    return _Graph()


def reset_default_graph():
    global _graph
    _graph = _Graph()


def get_default_graph():
    return _graph


def set_random_seed(seed):
    _graph.seed = seed


class _Ctx:
    """one Session.run: memo of evaluated nodes, the feeds, and the batches pulled from iterators"""

    def __init__(self, feed):
        self.feed = feed
        self.memo = {}
        self.pulled = {}


class Tensor:
    is_op = False

    def __init__(self, fn, inputs=(), name=None, dtype=None, shape=None):
        self._fn = fn
        self._inputs = tuple(inputs)
        self.dtype = dtype
        self._shape = shape          # callable -> list, or list, or None
        self.name = None
        if name is not None:
            self.name = _graph.unique(name)
            _graph.by_name[self.name] = self
            _graph.by_name[self.name + ":0"] = self

    # ---- evaluation
    def _eval(self, ctx):
        k = id(self)
        if k in ctx.memo:
            return ctx.memo[k]
        if k in ctx.feed:
            v = ctx.feed[k]
        else:
            v = self._fn(ctx, *[i._eval(ctx) if isinstance(i, Tensor) else i for i in self._inputs])
        ctx.memo[k] = v
        return v

    def eval(self, feed_dict=None, session=None):
        return (session or _default_session()).run(self, feed_dict=feed_dict)

    # ---- static shape (only what `x.shape.as_list()[1]` in the model files needs)
    @property
    def shape(self):
        s = self._shape
        if callable(s):
            s = s()
            self._shape = s
        return TensorShape(s)

    def get_shape(self):
        return self.shape

    # ---- python operators the model files use on tensors
    def __add__(self, o):
        return add(self, o)

    def __radd__(self, o):
        return add(o, self)

    def __sub__(self, o):
        return subtract(self, o)

    def __rsub__(self, o):
        return subtract(o, self)

    def __mul__(self, o):
        return multiply(self, o)

    def __rmul__(self, o):
        return multiply(o, self)

    def __truediv__(self, o):
        return _binary(lambda a, b: a / b, self, o)

    def __neg__(self):
        return _unary(lambda a: -a, self)

    def __getitem__(self, idx):
        def sl(ctx, a):
            return a[idx]

        def shp():
            try:
                base = self.shape.as_list()
            except ValueError:
                return None
            probe = np.empty([1 if d is None else d for d in base], dtype=np.bool_)[idx]
            out = list(probe.shape)
            # a full slice of an unknown leading dimension stays unknown
            ix = idx if isinstance(idx, tuple) else (idx,)
            j = 0
            for a, i in enumerate(ix):
                if isinstance(i, slice):
                    if base[a] is None and i == slice(None):
                        out[j] = None
                    j += 1
            return out

        return Tensor(sl, [self], dtype=self.dtype, shape=shp)

    def __repr__(self):
        return "<tf1-stand-in Tensor %s dtype=%s>" % (self.name or hex(id(self)), self.dtype)

    __hash__ = object.__hash__

    def __bool__(self):
        raise TypeError("a graph tensor has no truth value")


class Operation(Tensor):
    is_op = True

    def run(self, feed_dict=None, session=None):
        (session or _default_session()).run(self, feed_dict=feed_dict)


def _to_torch(v, dtype=None):
    if isinstance(v, torch.Tensor):
        t = v
    else:
        a = np.asarray(v)
        if a.dtype == np.float32 and F is torch.float64:
            a = a.astype(np.float64)          # exact
        t = torch.from_numpy(np.ascontiguousarray(a)) if a.shape else torch.tensor(a.item(), dtype=F if a.dtype.kind == 'f' else torch.int64)
    if dtype is not None:
        t = t.to(dtype.tt)
    elif t.is_floating_point() and t.dtype != F:
        t = t.to(F)
    elif t.dtype in (torch.int32, torch.int16, torch.int8, torch.uint8):
        t = t.to(torch.int64)
    return t


def _infer_dtype(t):
    return float32 if t.dtype.is_floating_point else int32


def convert_to_tensor(x, dtype=None, name=None):
    if isinstance(x, Tensor):
        return x
    dtype = _as_dtype(dtype)
    if isinstance(x, (list, tuple)) and any(isinstance(e, Tensor) for e in _flatten(x)):
        parts = [convert_to_tensor(e) for e in x]
        return Tensor(lambda ctx, *vs: torch.stack([v if isinstance(v, torch.Tensor) else _to_torch(v) for v in vs]),
                      parts, name=name, dtype=parts[0].dtype)
    t = _to_torch(x, dtype)
    return Tensor(lambda ctx: t, (), name=name, dtype=dtype or _infer_dtype(t), shape=list(t.shape))


def _flatten(x):
    for e in x:
        if isinstance(e, (list, tuple)):
            yield from _flatten(e)
        else:
            yield e


def constant(value, dtype=None, shape=None, name=None):
    return convert_to_tensor(value, dtype=dtype, name=name)


def _static(x):
    try:
        return x.shape.as_list()
    except ValueError:
        return None


# ------------------------------------------------------------------------------------------------ variables
class Variable(Tensor):
    def __init__(self, initial_value=None, trainable=True, name=None, dtype=None, **kw):
        if kw:
            raise TypeError("tf.Variable argument(s) %s not handled by the TF1 stand-in" % sorted(kw))
        dtype = _as_dtype(dtype)
        self._initial = initial_value if isinstance(initial_value, Tensor) else None
        self._initial_np = None if isinstance(initial_value, Tensor) else np.asarray(initial_value)
        if dtype is None:
            dtype = initial_value.dtype if isinstance(initial_value, Tensor) else (
                float32 if self._initial_np.dtype.kind == "f" else int32)
        self.trainable = bool(trainable) and dtype.is_float
        self.value = None
        super().__init__(None, (), name=name or "Variable", dtype=dtype,
                         shape=(lambda: _static(self._initial)) if self._initial is not None else list(
                             self._initial_np.shape))
        _graph.variables.append(self)

    def _eval(self, ctx):
        if self.value is None:
            raise RuntimeError("FailedPreconditionError: attempting to use uninitialized value " + self.name)
        return self.value

    def _initialise(self, ctx):
        if self._initial is not None:
            v = self._initial._eval(ctx).detach().clone()
        else:
            v = _to_torch(self._initial_np)
        if self.dtype.is_float:
            # TF stores the variable in float32: round the initial value through it, compute in float64
            v = v.to(torch.float32).to(F)
        else:
            v = v.to(torch.int64)
        if self.trainable:
            v.requires_grad_(True)
        self.value = v

    def numpy(self):
        return self.value.detach().numpy().copy()

    def load(self, array):
        v = _to_torch(np.asarray(array), self.dtype).clone()
        if self.trainable:
            v.requires_grad_(True)
        self.value = v

    __hash__ = object.__hash__


def global_variables():
    return list(_graph.variables)


def trainable_variables():
    return [v for v in _graph.variables if v.trainable]


def global_variables_initializer():
    g = _graph

    def init(ctx):
        for v in g.variables:
            v._initialise(ctx)
        for s in g.optimizer_slots:
            s()
        return None

    return Operation(init, (), name="init")


def local_variables_initializer():
    return Operation(lambda ctx: None, (), name="init_local")


def tables_initializer():
    return Operation(lambda ctx: None, (), name="init_all_tables")


def group(*ops, name=None):
    return Operation(lambda ctx, *vs: None, list(ops), name=name or "group")


class _VarScope:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        _graph.scope.append(self.name)
        return self

    def __exit__(self, *exc):
        _graph.scope.pop()
        return False

    def reuse_variables(self):
        return None


def variable_scope(name, **kw):
    return _VarScope(name)


def get_variable_scope():
    return _VarScope("")


class _NameScope:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        return self.name

    def __exit__(self, *exc):
        return False


def name_scope(name, *a, **kw):
    return _NameScope(name)


def get_variable(name, shape=None, dtype=None, initializer=None, trainable=True, **kw):
    if kw:
        raise TypeError("tf.get_variable argument(s) %s not handled by the TF1 stand-in" % sorted(kw))
    if initializer is None:
        initializer = _xavier_uniform()          # TF's default for float get_variable is glorot_uniform
    init = initializer(list(shape)) if callable(initializer) else initializer
    return Variable(init, trainable=trainable, name=name, dtype=dtype)


# ------------------------------------------------------------------------------------------------ random initial values
def _generator():
    g = _graph
    g.n_random += 1
    gen = torch.Generator()
    gen.manual_seed(((g.seed if g.seed is not None else 0) * 1000003 + g.n_random) % (2 ** 63))
    return gen


def _int_list(ctx, shape):
    out = []
    for d in shape:
        if isinstance(d, Tensor):
            d = d._eval(ctx)
        out.append(int(d))
    return out


def _shape_static(shape):
    return [None if isinstance(d, Tensor) else int(d) for d in shape]


def random_normal(shape, mean=0.0, stddev=1.0, dtype=None, seed=None, name=None):
    gen = _generator()
    return Tensor(lambda ctx: (torch.randn(_int_list(ctx, shape), generator=gen, dtype=torch.float64) * stddev + mean).to(F), (),
                  name=name, dtype=float32, shape=_shape_static(shape))


def random_uniform(shape, minval=0.0, maxval=1.0, dtype=None, seed=None, name=None):
    gen = _generator()
    return Tensor(lambda ctx: (torch.rand(_int_list(ctx, shape), generator=gen, dtype=torch.float64) * (maxval - minval) + minval).to(F),
                  (), name=name, dtype=float32, shape=_shape_static(shape))


def _xavier_uniform():
    def make(shape):
        fan_in, fan_out = (shape[0], shape[1]) if len(shape) == 2 else (shape[0], shape[0])
        lim = math.sqrt(6.0 / (fan_in + fan_out))
        return random_uniform(shape, -lim, lim)

    return make


def zeros(shape, dtype=float32, name=None):
    dtype = _as_dtype(dtype)
    return Tensor(lambda ctx: torch.zeros(_int_list(ctx, shape), dtype=dtype.tt), (), name=name, dtype=dtype,
                  shape=_shape_static(shape))


def ones(shape, dtype=float32, name=None):
    dtype = _as_dtype(dtype)
    return Tensor(lambda ctx: torch.ones(_int_list(ctx, shape), dtype=dtype.tt), (), name=name, dtype=dtype,
                  shape=_shape_static(shape))


# ------------------------------------------------------------------------------------------------ inputs
def placeholder(dtype, shape=None, name=None):
    dtype = _as_dtype(dtype)

    def missing(ctx):
        raise RuntimeError("InvalidArgumentError: you must feed a value for placeholder tensor " + str(t.name))

    t = Tensor(missing, (), name=name or "Placeholder", dtype=dtype, shape=None if shape is None else list(shape))
    t._is_placeholder = True
    return t


def placeholder_with_default(input, shape, name=None):
    src = convert_to_tensor(input)
    t = Tensor(lambda ctx, v: v, [src], name=name or "PlaceholderWithDefault", dtype=src.dtype,
               shape=None if shape is None else list(shape))
    t._is_placeholder = True
    return t


class errors:
    class OutOfRangeError(Exception):
        pass


class BatchSource:
    """Stands in for `tf.data ... make_one_shot_iterator().get_next()`: `tensors` is the dict of component
    tensors; every Session.run that needs any component pulls ONE element (all components from the same
    element) and raises OutOfRangeError at the end, as an iterator's get_next op does."""

    def __init__(self, batches, spec):
        self._it = iter(batches)
        self.n_pulled = 0
        self.tensors = {}
        for key, (dtype, shape) in spec.items():
            self.tensors[key] = Tensor(self._component(key, _as_dtype(dtype)), (), name="IteratorGetNext_" + key,
                                       dtype=_as_dtype(dtype), shape=list(shape))

    def _component(self, key, dtype):
        def pull(ctx):
            k = id(self)
            if k not in ctx.pulled:
                try:
                    ctx.pulled[k] = next(self._it)
                except StopIteration:
                    raise errors.OutOfRangeError("End of sequence") from None
                self.n_pulled += 1
            return _to_torch(ctx.pulled[k][key], dtype)

        return pull


# ------------------------------------------------------------------------------------------------ ops
def _bshape(a, b):
    sa, sb = _static(a), _static(b)
    if sa is None or sb is None:
        return None
    n = max(len(sa), len(sb))
    sa, sb = [1] * (n - len(sa)) + sa, [1] * (n - len(sb)) + sb
    out = []
    for x, y in zip(sa, sb):
        if x is None or y is None:
            out.append(None if (x in (None, 1) and y in (None, 1)) else (y if x is None else x))
        else:
            out.append(max(x, y))
    return out


def _binary(f, a, b, name=None):
    a, b = convert_to_tensor(a), convert_to_tensor(b)

    def run(ctx, x, y):
        if x.dtype != y.dtype:
            if x.dtype.is_floating_point != y.dtype.is_floating_point and x.dim() and y.dim():
                raise TypeError("TypeError: mixed float/int operands (TF does not promote): %s" % name)
        return f(x, y)

    return Tensor(run, [a, b], name=name, dtype=a.dtype if a.dtype and a.dtype.is_float else b.dtype,
                  shape=lambda: _bshape(a, b))


def _unary(f, a, name=None, dtype=None):
    a = convert_to_tensor(a)
    return Tensor(lambda ctx, x: f(x), [a], name=name, dtype=dtype or a.dtype, shape=lambda: _static(a))


def add(x, y, name=None):
    return _binary(lambda a, b: a + b, x, y, name)


def subtract(x, y, name=None):
    return _binary(lambda a, b: a - b, x, y, name)


def multiply(x, y, name=None):
    return _binary(lambda a, b: a * b, x, y, name)


def div_no_nan(x, y, name=None):
    def f(a, b):
        nz = b != 0
        return torch.where(nz, a / torch.where(nz, b, torch.ones_like(b)), torch.zeros_like(a * b))

    return _binary(f, x, y, name)


def square(x, name=None):
    return _unary(lambda a: a * a, x, name)


def identity(x, name=None):
    return _unary(lambda a: a, x, name)


def _axes(axis):
    if axis is None:
        return None
    return tuple(axis) if isinstance(axis, (list, tuple)) else (int(axis),)


def _reduce_shape(x, axis, keep):
    def f():
        s = _static(x)
        if s is None:
            return None
        ax = _axes(axis)
        if ax is None:
            return [1] * len(s) if keep else []
        ax = [a % len(s) for a in ax]
        return [1 if i in ax else d for i, d in enumerate(s)] if keep else [d for i, d in enumerate(s) if i not in ax]

    return f


def _keep(keepdims, keep_dims):
    return bool(keepdims) or bool(keep_dims)


def reduce_sum(input_tensor, axis=None, keepdims=None, name=None, keep_dims=None, reduction_indices=None):
    x = convert_to_tensor(input_tensor)
    keep, ax = _keep(keepdims, keep_dims), _axes(axis if axis is not None else reduction_indices)
    return Tensor(lambda ctx, a: a.sum() if ax is None and not keep else a.sum(dim=ax if ax is not None else tuple(
        range(a.dim())), keepdim=keep), [x], name=name, dtype=x.dtype, shape=_reduce_shape(x, ax, keep))


def reduce_mean(input_tensor, axis=None, keepdims=None, name=None, keep_dims=None):
    x = convert_to_tensor(input_tensor)
    keep, ax = _keep(keepdims, keep_dims), _axes(axis)
    return Tensor(lambda ctx, a: a.mean() if ax is None and not keep else a.mean(dim=ax if ax is not None else tuple(
        range(a.dim())), keepdim=keep), [x], name=name, dtype=x.dtype, shape=_reduce_shape(x, ax, keep))


def count_nonzero(input_tensor, axis=None, keepdims=None, dtype=int64, name=None, keep_dims=None):
    x = convert_to_tensor(input_tensor)
    keep, ax, dt = _keep(keepdims, keep_dims), _axes(axis), _as_dtype(dtype)
    return Tensor(lambda ctx, a: (a.detach() != 0).to(dt.tt).sum(dim=ax if ax is not None else tuple(range(a.dim())),
                                                                keepdim=keep), [x], name=name, dtype=dt,
                  shape=_reduce_shape(x, ax, keep))


def reshape(tensor, shape, name=None):
    if _any_eager(tensor, shape):
        a, shp = _eager_value(tensor), [int(d) for d in _eager_value(shape).reshape(-1)]
        if int(np.prod(shp)) != a.size:
            raise ValueError("InvalidArgumentError: Input to reshape is a tensor with %d values, but the requested shape "
                             "has %d" % (a.size, int(np.prod(shp))))
        return _Eager(a.reshape(shp))
    x = convert_to_tensor(tensor)

    def run(ctx, a):
        return a.reshape(_int_list(ctx, shape))

    def shp():
        tgt = _shape_static(shape)
        if -1 in tgt:
            i = tgt.index(-1)
            s = _static(x)
            if s is not None and None not in s and None not in tgt:
                rest = int(np.prod([d for j, d in enumerate(tgt) if j != i])) or 1
                tgt[i] = int(np.prod(s)) // rest
            else:
                tgt[i] = None
        return tgt

    return Tensor(run, [x], name=name, dtype=x.dtype, shape=shp)


def concat(values, axis, name=None):
    vs = [convert_to_tensor(v) for v in values]

    def run(ctx, *xs):
        kinds = {x.dtype.is_floating_point for x in xs}
        if len(kinds) != 1:
            raise TypeError("concat of float and integer tensors")
        return torch.cat(list(xs), dim=axis)

    def shp():
        ss = [_static(v) for v in vs]
        if any(s is None for s in ss):
            return None
        out = list(ss[0])
        a = axis % len(out)
        for j in range(len(out)):
            col = [s[j] for s in ss]
            if j == a:
                out[j] = None if None in col else sum(col)
            else:
                known = [c for c in col if c is not None]
                out[j] = known[0] if known else None
        return out

    return Tensor(run, vs, name=name, dtype=vs[0].dtype, shape=shp)


def tile(input, multiples, name=None):
    x = convert_to_tensor(input)

    def shp():
        s = _static(x)
        if s is None:
            return None
        return [None if (isinstance(m, Tensor) or d is None) else d * int(m) for d, m in zip(s, multiples)]

    return Tensor(lambda ctx, a: a.repeat(_int_list(ctx, multiples)), [x], name=name, dtype=x.dtype, shape=shp)


def matmul(a, b, name=None):
    a, b = convert_to_tensor(a), convert_to_tensor(b)

    def shp():
        sa, sb = _static(a), _static(b)
        return None if sa is None or sb is None else [sa[0], sb[1]]

    return Tensor(lambda ctx, x, y: x @ y, [a, b], name=name, dtype=a.dtype, shape=shp)


def _matmul_t(a, b, transpose_a=False, transpose_b=False, name=None):
    """tf.matmul incl. the batched form with transpose flags (test_model/DCN.py:85: [B,D,1] x [B,D,1]^T -> [B,D,D])"""
    if not transpose_a and not transpose_b:
        return matmul(a, b, name=name)
    a, b = convert_to_tensor(a), convert_to_tensor(b)

    def run(ctx, x, y):
        return (x.transpose(-1, -2) if transpose_a else x) @ (y.transpose(-1, -2) if transpose_b else y)
    return Tensor(run, [a, b], name=name, dtype=a.dtype, shape=None)


def tensordot(a, b, axes, name=None):
    """tf.tensordot with an integer `axes`: contracts the last `axes` axes of a with the first `axes` of b"""
    if not isinstance(axes, int):
        raise NotImplementedError("tensordot with explicit axis lists")
    a, b = convert_to_tensor(a), convert_to_tensor(b)
    return Tensor(lambda ctx, x, y: torch.tensordot(x, y, dims=axes), [a, b], name=name, dtype=a.dtype, shape=None)


def sigmoid(x, name=None):
    return _unary(torch.sigmoid, x, name or "Sigmoid")          # TF names the op: test_model/DCN.py:190 fetches "Sigmoid:0"


def _sigmoid_cross_entropy_with_logits(_sentinel=None, labels=None, logits=None, name=None):
    """documented formula: max(x, 0) - x * z + log(1 + exp(-|x|))"""
    if _sentinel is not None:
        raise ValueError("Only call `sigmoid_cross_entropy_with_logits` with named arguments (labels=..., logits=...)")
    z, x = convert_to_tensor(labels), convert_to_tensor(logits)
    return Tensor(lambda ctx, zz, xx: torch.clamp(xx, min=0) - xx * zz + torch.log1p(torch.exp(-xx.abs())), [z, x], name=name,
                  dtype=x.dtype, shape=lambda: _static(x))


def _relu(x, name=None):
    return _unary(torch.relu, x, name)


def _dropout(x, keep_prob=None, noise_shape=None, seed=None, name=None, rate=None):
    if rate is not None:
        keep_prob = 1 - rate
    if isinstance(keep_prob, Tensor) or float(keep_prob) != 1.0:
        raise NotImplementedError("dropout with keep_prob != 1 draws from TF's generator: not reproducible here")
    return convert_to_tensor(x)                      # tf.nn.dropout returns x itself when keep_prob is 1


def _embedding_lookup(params, ids, partition_strategy="mod", name=None, validate_indices=True, max_norm=None):
    if isinstance(params, (list, tuple)):
        if len(params) != 1:
            raise NotImplementedError("partitioned embedding_lookup")
        params = params[0]
    p, i = convert_to_tensor(params), convert_to_tensor(ids)

    def run(ctx, table, idx):
        if idx.dtype.is_floating_point:
            raise TypeError("embedding_lookup ids must be integers")
        if idx.numel() and (int(idx.min()) < 0 or int(idx.max()) >= table.shape[0]):
            raise IndexError("InvalidArgumentError: indices[...] = %d is not in [0, %d)" % (
                int(idx.max()) if int(idx.max()) >= table.shape[0] else int(idx.min()), table.shape[0]))
        return table[idx]

    def shp():
        sp, si = _static(p), _static(i)
        return None if sp is None or si is None else si + sp[1:]

    t = Tensor(run, [p, i], name=name, dtype=p.dtype, shape=shp)
    t._is_gather = True          # its gradient w.r.t. `params` is an IndexedSlices (sparse optimizer applies look at this)
    return t


nn = types.SimpleNamespace(embedding_lookup=_embedding_lookup, dropout=_dropout, relu=_relu, sigmoid=sigmoid,
                           sigmoid_cross_entropy_with_logits=_sigmoid_cross_entropy_with_logits)


# ------------------------------------------------------------------------------------------------ losses / regularisers
def _log_loss(labels, predictions, weights=1.0, epsilon=1e-7, scope=None):
    if weights != 1.0:
        raise NotImplementedError("weighted log_loss")
    y, p = convert_to_tensor(labels), convert_to_tensor(predictions)

    def run(ctx, yy, pp):
        if yy.shape != pp.shape:
            raise ValueError("Shapes %s and %s are incompatible" % (tuple(pp.shape), tuple(yy.shape)))
        yy = yy.to(F)
        per = -yy * torch.log(pp + epsilon) - (1 - yy) * torch.log(1 - pp + epsilon)
        return per.sum() / per.numel()               # Reduction.SUM_BY_NONZERO_WEIGHTS with unit weights

    return Tensor(run, [y, p], dtype=float32, shape=[])


losses = types.SimpleNamespace(log_loss=_log_loss)


def _l2_regularizer(scale, scope=None):
    def reg(w):
        return _unary(lambda a: scale * (a * a).sum() / 2, w)        # scale * tf.nn.l2_loss(w)

    return reg


def _l1_regularizer(scale, scope=None):
    def reg(w):
        return _unary(lambda a: scale * a.abs().sum(), w)

    return reg


contrib = types.SimpleNamespace(layers=types.SimpleNamespace(
    l2_regularizer=_l2_regularizer, l1_regularizer=_l1_regularizer, xavier_initializer=lambda *a, **k: _xavier_uniform()))


# ------------------------------------------------------------------------------------------------ training
def _exponential_decay(learning_rate, global_step, decay_steps, decay_rate, staircase=False, name=None):
    gs = convert_to_tensor(global_step)

    def run(ctx, step):
        e = float(step) / float(decay_steps)
        if staircase:
            e = math.floor(e)
        return torch.tensor(float(learning_rate) * float(decay_rate) ** e, dtype=F)

    return Tensor(run, [gs], name=name, dtype=float32, shape=[])


class _AdamOptimizer:
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-08, use_locking=False, name="Adam"):
        self.lr, self.b1, self.b2, self.eps = learning_rate, beta1, beta2, epsilon
        self.m, self.v = {}, {}
        self.b1_power, self.b2_power = None, None

    def _reset(self):
        self.m, self.v = {}, {}
        self.b1_power, self.b2_power = self.b1, self.b2

    def minimize(self, loss, global_step=None, var_list=None):
        g = _graph
        g.optimizer_slots.append(self._reset)            # slots are global variables: (re)set by the initializer
        self._reset()
        lr = convert_to_tensor(self.lr)

        def run(ctx):
            vs = list(var_list) if var_list is not None else [v for v in g.variables if v.trainable]
            lv = loss._eval(ctx)
            grads = torch.autograd.grad(lv, [v.value for v in vs], allow_unused=True, retain_graph=True)
            if all(gr is None for gr in grads):
                raise ValueError("No gradients provided for any variable")
            lr_t = float(lr._eval(ctx)) * math.sqrt(1 - self.b2_power) / (1 - self.b1_power)
            with torch.no_grad():
                for var, gr in zip(vs, grads):
                    if gr is None:
                        continue
                    k = id(var)
                    if k not in self.m:
                        self.m[k], self.v[k] = torch.zeros_like(var.value), torch.zeros_like(var.value)
                    self.m[k].mul_(self.b1).add_(gr, alpha=1 - self.b1)
                    self.v[k].mul_(self.b2).add_(gr * gr, alpha=1 - self.b2)
                    new = var.value - lr_t * self.m[k] / (self.v[k].sqrt() + self.eps)
                    var.value = new.detach().requires_grad_(True)
                self.b1_power *= self.b1
                self.b2_power *= self.b2
                if global_step is not None:
                    global_step.value = global_step.value + 1
            return None

        return Operation(run, (), name="Adam")


def _consumers(root):
    """id(tensor) -> nodes of the graph under `root` that read it"""
    seen, out, stack = set(), {}, [root]
    while stack:
        n = stack.pop()
        if id(n) in seen:
            continue
        seen.add(id(n))
        for i in getattr(n, "_inputs", ()):
            if isinstance(i, Tensor):
                out.setdefault(id(i), []).append(n)
                stack.append(i)
    return out


class _FtrlOptimizer:
    """tf.train.FtrlOptimizer (ftrl.py -> ApplyFtrl / SparseApplyFtrl, published kernels; l2_shrinkage 0):
        accum' = accum + g^2 ; linear += g - (accum'^(-p) - accum^(-p)) / lr * var ; quadratic = accum'^(-p) / lr + 2 l2
        var = (sign(linear) l1 - linear) / quadratic if |linear| > l1 else 0          (p = learning_rate_power)
    A variable read ONLY through embedding_lookup has an IndexedSlices gradient: the sparse kernel touches the looked-up
    rows alone (duplicates summed first); any other variable takes the dense kernel."""

    def __init__(self, learning_rate, learning_rate_power=-0.5, initial_accumulator_value=0.1, l1_regularization_strength=0.0,
                 l2_regularization_strength=0.0, use_locking=False, name="Ftrl", **kw):
        if kw:
            raise NotImplementedError("FtrlOptimizer arguments %s" % sorted(kw))
        self.lr, self.p, self.a0 = learning_rate, float(learning_rate_power), float(initial_accumulator_value)
        self.l1, self.l2 = float(l1_regularization_strength), float(l2_regularization_strength)
        self.accum, self.linear = {}, {}

    def _reset(self):
        self.accum, self.linear = {}, {}

    def minimize(self, loss, global_step=None, var_list=None):
        g = _graph
        g.optimizer_slots.append(self._reset)
        self._reset()
        lr = convert_to_tensor(self.lr)
        users = _consumers(loss)

        def run(ctx):
            vs = list(var_list) if var_list is not None else [v for v in g.variables if v.trainable]
            lv = loss._eval(ctx)
            grads = torch.autograd.grad(lv, [v.value for v in vs], allow_unused=True, retain_graph=True)
            if all(gr is None for gr in grads):
                raise ValueError("No gradients provided for any variable")
            lrv = float(lr._eval(ctx))
            with torch.no_grad():
                for var, gr in zip(vs, grads):
                    if gr is None:
                        continue
                    k = id(var)
                    if k not in self.accum:
                        self.accum[k], self.linear[k] = torch.full_like(var.value, self.a0), torch.zeros_like(var.value)
                    acc, lin, w = self.accum[k], self.linear[k], var.value
                    acc_new = acc + gr * gr
                    lin_new = lin + gr - (acc_new.pow(-self.p) - acc.pow(-self.p)) / lrv * w
                    quad = acc_new.pow(-self.p) / lrv + 2 * self.l2
                    w_new = torch.where(lin_new.abs() > self.l1, (torch.sign(lin_new) * self.l1 - lin_new) / quad, torch.zeros_like(w))
                    use = users.get(k, [])
                    if use and all(getattr(n, "_is_gather", False) and n._inputs[0] is var for n in use):
                        rows = torch.zeros(w.shape[0], dtype=torch.bool)
                        for n in use:
                            rows[n._inputs[1]._eval(ctx).reshape(-1)] = True
                        mk = rows.reshape(-1, *([1] * (w.dim() - 1)))
                        acc_new, lin_new, w_new = torch.where(mk, acc_new, acc), torch.where(mk, lin_new, lin), torch.where(mk, w_new, w)
                    self.accum[k], self.linear[k] = acc_new, lin_new
                    var.value = w_new.detach().requires_grad_(True)
                if global_step is not None:
                    global_step.value = global_step.value + 1
            return None

        return Operation(run, (), name="Ftrl")


class _Saver:
    def __init__(self, *a, **k):
        pass

    def save(self, sess, save_path, global_step=None, **k):
        return save_path

    def restore(self, sess, save_path):
        raise NotImplementedError("checkpoint restore is outside the path the stand-in runs")


train = types.SimpleNamespace(AdamOptimizer=_AdamOptimizer, FtrlOptimizer=_FtrlOptimizer, exponential_decay=_exponential_decay, Saver=_Saver,
                              latest_checkpoint=lambda d: None)


# ------------------------------------------------------------------------------------------------ session
_sessions = []


def _default_session():
    if not _sessions:
        raise ValueError("Cannot evaluate tensor using `eval()`: No default session is registered")
    return _sessions[-1]


class Session:
    def __init__(self, target="", graph=None, config=None):
        self.graph = graph if graph is not None else _graph      # a fresh Graph() until loader.load binds a saved one

    def __enter__(self):
        _sessions.append(self)
        return self

    def __exit__(self, *exc):
        _sessions.pop()
        return False

    def close(self):
        pass

    def _lookup(self, f):
        if isinstance(f, str):
            if f not in self.graph.by_name:
                raise KeyError("The name %r refers to a Tensor which does not exist" % f)
            return self.graph.by_name[f]
        return f

    def run(self, fetches, feed_dict=None):
        feed = {}
        for k, v in (feed_dict or {}).items():
            t = self._lookup(k)
            val = _to_torch(np.asarray(v), t.dtype)
            s = t._shape if isinstance(t._shape, list) else None
            if s is not None and (len(s) != val.dim() or any(d is not None and d != n for d, n in zip(s, val.shape))):
                raise ValueError("Cannot feed value of shape %s for Tensor %r, which has shape %s" % (
                    tuple(val.shape), t.name, s))
            feed[id(t)] = val
        ctx = _Ctx(feed)
        single = not isinstance(fetches, (list, tuple))
        fl = [self._lookup(f) for f in ([fetches] if single else fetches)]
        out = [None] * len(fl)
        # values first (they see the variables as they were), then the state-changing ops
        for j, f in enumerate(fl):
            if not f.is_op:
                v = convert_to_tensor(f)._eval(ctx)
                a = v.detach().numpy().copy()
                if a.dtype.kind == "f" and f.dtype is not float64:
                    pass                           # kept in float64: the fixtures hold the un-rounded values
                out[j] = a if a.shape else a[()]
        for j, f in enumerate(fl):
            if f.is_op:
                f._eval(ctx)
        return out[0] if single else out


# ------------------------------------------------------------------------------------------------ tf.data + string ops
# `Dataset.map` functions of the reference (utils/data_loader_load.py:187-240, utils/data_loader_textline.py:4-36) run
# once per ELEMENT here, on eager values: numpy arrays (dtype=object for strings) behind a thin wrapper.
class _Eager:
    def __init__(self, a):
        self.a = a if isinstance(a, np.ndarray) else np.asarray(a)

    def __getitem__(self, idx):
        return _Eager(self.a[idx])

    @property
    def shape(self):
        return TensorShape(list(self.a.shape))


def _eager_value(x):
    if isinstance(x, _Eager):
        return x.a
    if isinstance(x, (list, tuple)):
        return np.array([_eager_value(e) for e in x], dtype=object if any(
            isinstance(e, _Eager) and e.a.dtype == object for e in x) else None)
    return np.asarray(x)


def _any_eager(*xs):
    for x in xs:
        if isinstance(x, _Eager) or (isinstance(x, (list, tuple)) and any(isinstance(e, _Eager) for e in x)):
            return True
    return False


class _SparseStrings:
    def __init__(self, indices, values, dense_shape):
        self.indices, self.values, self.dense_shape = _Eager(indices), _Eager(values), _Eager(dense_shape)


def string_split(source, delimiter=" ", skip_empty=True):
    """tf.string_split: the tokens of every entry of a 1-D string tensor, row-major; empty tokens are skipped by default;
    dense_shape = [entries, longest row]"""
    src = _eager_value(source).reshape(-1)
    idx, vals, longest = [], [], 0
    for r, entry in enumerate(src):
        toks = str(entry).split(delimiter) if delimiter != "" else list(str(entry))
        if skip_empty:
            toks = [t for t in toks if t != ""]
        for c, t in enumerate(toks):
            idx.append((r, c))
            vals.append(t)
        longest = max(longest, len(toks))
    v = np.empty(len(vals), dtype=object)
    v[:] = vals
    return _SparseStrings(np.asarray(idx, dtype=np.int64).reshape(-1, 2), v, np.asarray([len(src), longest], dtype=np.int64))


def string_to_number(string_tensor, out_type=float32, name=None):
    """tf.string_to_number: every entry must be a valid number of the output type (InvalidArgumentError otherwise);
    float32 results are the float32 nearest to the decimal (Python float() then a float32 cast: the same value except
    for decimals within 2^-53 of a float32 rounding boundary)"""
    a = _eager_value(string_tensor)
    dt = _as_dtype(out_type)
    flat = a.reshape(-1)
    if dt.is_float:
        out = np.empty(flat.shape, dtype=np.float32)
        for i, t in enumerate(flat):
            t = str(t)
            if t != t.strip() or "_" in t or "x" in t.lower() and "inf" not in t.lower():
                raise ValueError("InvalidArgumentError: StringToNumberOp could not correctly convert string: " + t)
            try:
                out[i] = np.float32(float(t))
            except ValueError:
                raise ValueError("InvalidArgumentError: StringToNumberOp could not correctly convert string: " + t) from None
    else:
        out = np.empty(flat.shape, dtype=np.int32 if out_type is int32 else np.int64)
        for i, t in enumerate(flat):
            t = str(t)
            ok = t.lstrip("+-").isdigit() and t == t.strip() and len(t) - len(t.lstrip("+-")) <= 1 and t.isascii()
            if not ok or not (np.iinfo(out.dtype).min <= int(t) <= np.iinfo(out.dtype).max):
                raise ValueError("InvalidArgumentError: StringToNumberOp could not correctly convert string: " + t)
            out[i] = int(t)
    return _Eager(out.reshape(a.shape))


def split(value, num_or_size_splits, axis=0, name=None):
    a = _eager_value(value)
    if not isinstance(num_or_size_splits, int) or a.shape[axis] % num_or_size_splits:
        raise ValueError("InvalidArgumentError: dimension %d is not divisible by %r" % (a.shape[axis], num_or_size_splits))
    return [_Eager(p) for p in np.split(a, num_or_size_splits, axis=axis)]


def squeeze(input, axis=None, name=None):
    a = _eager_value(input)
    ax = _axes(axis)
    if ax is not None and any(a.shape[i] != 1 for i in ax):
        raise ValueError("InvalidArgumentError: can not squeeze a dimension that is not 1")
    return _Eager(np.squeeze(a, axis=ax))


class _Dataset:
    def __init__(self, make):
        self._make = make                     # () -> iterator over elements

    def map(self, fn, num_parallel_calls=None):
        def to_np(v):
            return v.a if isinstance(v, _Eager) else np.asarray(v, dtype=np.int32 if isinstance(v, int) else None)

        def gen():
            for e in self._make():
                out = fn(_Eager(np.array(e, dtype=object)))
                yield {k: to_np(v) for k, v in out.items()}
        return _Dataset(gen)

    def batch(self, batch_size, drop_remainder=False):
        def gen():
            buf = []
            for e in self._make():
                buf.append(e)
                if len(buf) == batch_size:
                    yield self._stack(buf)
                    buf = []
            if buf and not drop_remainder:
                yield self._stack(buf)
        return _Dataset(gen)

    @staticmethod
    def _stack(buf):
        out = {}
        for k in buf[0]:
            shapes = {e[k].shape for e in buf}
            if len(shapes) != 1:
                raise ValueError("InvalidArgumentError: Cannot batch tensors with different shapes in component %s: %s" % (
                    k, sorted(shapes)))
            out[k] = np.stack([e[k] for e in buf])
        return out

    def repeat(self, count=None):
        def gen():
            n = 0
            while count is None or n < count:
                yield from self._make()
                n += 1
        return _Dataset(gen)

    def prefetch(self, n):
        return self

    def make_one_shot_iterator(self):
        return _OneShotIterator(self._make())


class _OneShotIterator:
    def __init__(self, it):
        self._it = it

    def get_next(self):
        import itertools
        try:
            first = next(self._it)
        except StopIteration:
            raise ValueError("the stand-in needs one element to learn the component types") from None
        spec = {k: (float32 if v.dtype.kind == "f" else int64, [None] + list(v.shape[1:])) for k, v in first.items()}
        return BatchSource(itertools.chain([first], self._it), spec).tensors


def _text_line_dataset(filenames, **kw):
    """tf.data.TextLineDataset: the lines of the files in order, line terminator removed"""
    if kw:
        raise TypeError("TextLineDataset argument(s) %s not handled by the TF1 stand-in" % sorted(kw))
    names = [filenames] if isinstance(filenames, str) else list(filenames)

    def gen():
        for fn in names:
            with open(fn, "r", newline="\n") as f:
                for line in f:
                    yield line[:-1] if line.endswith("\n") else line
    return _Dataset(gen)


data = types.SimpleNamespace(TextLineDataset=_text_line_dataset)


# ------------------------------------------------------------------------------------------------ file system
def _fast_gfile(name, mode="r"):
    return open(name, mode)                      # tf.gfile.FastGFile on a local path: a Python file object


def _list_directory(path):
    import os
    return os.listdir(path)                      # tf.gfile.ListDirectory: entry names, order unspecified


gfile = types.SimpleNamespace(FastGFile=_fast_gfile, ListDirectory=_list_directory)


# ------------------------------------------------------------------------------------------------ saved-model stubs
_saved_models = {}


class _SavedModelBuilder:
    """SavedModelBuilder(dir).add_meta_graph_and_variables(sess, tags, ...).save(): here the "file" is an entry of a
    process-wide registry holding the graph (its names and variables), which is all `loader.load` + `run("score:0")` need"""

    def __init__(self, export_dir):
        if export_dir in _saved_models:
            raise AssertionError("Export directory already exists, and isn't empty: " + str(export_dir))
        self._dir, self._graph = export_dir, None

    def add_meta_graph_and_variables(self, sess, tags, signature_def_map=None, clear_devices=False, legacy_init_op=None,
                                     **kw):
        if kw:
            raise TypeError("SavedModelBuilder argument(s) %s not handled by the TF1 stand-in" % sorted(kw))
        for v in sess.graph.variables:
            if v.value is None:
                raise RuntimeError("FailedPreconditionError: attempting to save uninitialized value " + v.name)
        self._graph, self._tags = sess.graph, list(tags)

    def save(self):
        if self._graph is None:
            raise AssertionError("Graph state including variables and assets has not been saved yet")
        _saved_models[self._dir] = (self._graph, self._tags)
        return self._dir


def _loader_load(sess, tags, export_dir):
    if export_dir not in _saved_models:
        raise IOError("SavedModel file does not exist at: " + str(export_dir))
    graph, saved_tags = _saved_models[export_dir]
    if list(tags) != saved_tags:
        raise RuntimeError("MetaGraphDef associated with tags %r could not be found in SavedModel" % (list(tags),))
    sess.graph = graph


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    return m


def install():
    """register this file as `tensorflow` (and the few submodules the model files import)"""
    me = sys.modules[__name__]
    tf = _module("tensorflow")
    for k in ("float32", "float64", "int32", "int64", "Graph", "reset_default_graph", "get_default_graph",
              "set_random_seed", "Tensor", "Variable", "constant", "convert_to_tensor", "global_variables",
              "trainable_variables", "global_variables_initializer", "local_variables_initializer",
              "tables_initializer", "group", "variable_scope", "get_variable_scope", "name_scope", "get_variable",
              "random_normal", "random_uniform", "zeros", "ones", "placeholder", "placeholder_with_default",
              "errors", "add", "subtract", "multiply", "div_no_nan", "square", "identity", "reduce_sum",
              "reduce_mean", "count_nonzero", "reshape", "concat", "tile", "tensordot", "sigmoid", "nn", "losses",
              "contrib", "train", "Session", "gfile", "data", "string_split", "string_to_number", "split", "squeeze"):
        setattr(tf, k, getattr(me, k))
    tf.matmul = _matmul_t
    tag_constants = _module("tensorflow.python.saved_model.tag_constants", SERVING="serve")
    signature_constants = _module("tensorflow.python.saved_model.signature_constants",
                                  PREDICT_METHOD_NAME="tensorflow/serving/predict",
                                  DEFAULT_SERVING_SIGNATURE_DEF_KEY="serving_default")
    signature_def_utils = _module("tensorflow.python.saved_model.signature_def_utils",
                                  build_signature_def=lambda inputs=None, outputs=None, method_name=None: dict(
                                      inputs=inputs, outputs=outputs, method_name=method_name))
    utils = _module("tensorflow.python.saved_model.utils", build_tensor_info=lambda t: t)
    builder = _module("tensorflow.python.saved_model.builder", SavedModelBuilder=_SavedModelBuilder)
    sm = _module("tensorflow.python.saved_model", builder=builder, signature_constants=signature_constants,
                 signature_def_utils=signature_def_utils, tag_constants=tag_constants, utils=utils)
    py = _module("tensorflow.python", saved_model=sm)
    tf.python = py
    tf.saved_model = types.SimpleNamespace(tag_constants=tag_constants, loader=types.SimpleNamespace(load=_loader_load))
    mods = {"tensorflow": tf, "tensorflow.python": py, "tensorflow.python.saved_model": sm,
            "tensorflow.python.saved_model.builder": builder,
            "tensorflow.python.saved_model.signature_constants": signature_constants,
            "tensorflow.python.saved_model.signature_def_utils": signature_def_utils,
            "tensorflow.python.saved_model.tag_constants": tag_constants,
            "tensorflow.python.saved_model.utils": utils}
    for name, m in mods.items():
        if name in sys.modules and getattr(sys.modules[name], "__file__", None):
            raise RuntimeError("a real %s is importable: use it instead of the stand-in" % name)
        sys.modules[name] = m
    return tf
