"""`tensorflow` stand-in used ONLY to execute the reference's own Python source
(/root/reference/KGE) in this container — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

TensorFlow 2.5 is not installable here.  This package implements, on
torch-CPU fp32, exactly the ~50 `tf.*` entry points the reference touches, with the TF
semantics of SURVEY.md Appendix B:
  * `tf.nn.embedding_lookup` on a Variable under a `GradientTape` yields IndexedSlices
    gradients; several lookups of one variable are concatenated, not merged (B.1);
  * `tf.clip_by_norm` on IndexedSlices takes the norm over the un-merged values (B.3);
  * `tf.optimizers.Adam` sparse apply = dedup-sum, then DENSE m/v decay and DENSE
    variable move (B.2);
  * `tf.clip_by_value` passes gradient on the closed interval (B.4).
It is used by tests/golden/make_golden.py to generate the committed golden vectors and
never imported by the product package.  It is NOT TensorFlow: results pin the
reference's *source formulas*, not TF's kernels ("parity unpinned" at that boundary).
"""
import math as _math
import types as _types

import numpy as _np
import torch as _torch

__version__ = "2.5.0-shim"

float32 = _torch.float32
float64 = _torch.float64
int16 = _torch.int16
int32 = _torch.int32
int64 = _torch.int64
bool = _torch.bool  # noqa: A001
complex64 = _torch.complex64

_NP2T = {_np.dtype("float32"): _torch.float32, _np.dtype("float64"): _torch.float64,
         _np.dtype("int32"): _torch.int32, _np.dtype("int64"): _torch.int64,
         _np.dtype("int16"): _torch.int16}


def _dt(d):
    if d is None or isinstance(d, _torch.dtype):
        return d
    return _NP2T[_np.dtype(d)]


class Tensor(_torch.Tensor):
    """torch.Tensor with the few tf.Tensor conveniences the reference uses."""

    def numpy(self):
        t = self.detach().as_subclass(_torch.Tensor)
        a = t.numpy()
        return a[()] if a.ndim == 0 else a

    def __len__(self):
        return self.shape[0]

    def __array__(self, dtype=None, copy=None):
        a = self.detach().as_subclass(_torch.Tensor).numpy()
        return a.astype(dtype) if dtype is not None else a

    def __hash__(self):
        return id(self)


def _t(x, dtype=None):
    """anything -> Tensor (no copy when already a torch tensor)."""
    if isinstance(x, IndexedSlices):
        raise TypeError("IndexedSlices used as a dense tensor")
    if isinstance(x, _torch.Tensor):
        out = x
    else:
        a = _np.asarray(x)
        if a.dtype == _np.float64 and dtype is None and not isinstance(x, _np.ndarray):
            a = a.astype(_np.float32)          # python floats are float32 constants in TF
        out = _torch.as_tensor(a)
    if dtype is not None and out.dtype != _dt(dtype):
        out = out.to(_dt(dtype))
    return out.as_subclass(Tensor)


_ACTIVE_TAPES = []


class Variable(Tensor):
    """tf.Variable: a leaf that requires grad; `assign` writes in place."""

    @staticmethod
    def __new__(cls, initial_value, name=None, dtype=None, trainable=True):
        v = _t(initial_value, dtype).detach().clone().as_subclass(_torch.Tensor)
        if v.dtype == _torch.float64:
            v = v.float()
        obj = _torch.Tensor._make_subclass(cls, v, True)
        obj._tf_name = name
        return obj

    def __init__(self, *a, **k):
        pass

    def assign(self, value):
        with _torch.no_grad():
            self.as_subclass(_torch.Tensor).copy_(_t(value).as_subclass(_torch.Tensor))
        return self

    def assign_sub(self, value):
        with _torch.no_grad():
            self.as_subclass(_torch.Tensor).sub_(_t(value).as_subclass(_torch.Tensor))
        return self

    def value(self):
        return self

    def __deepcopy__(self, memo):
        return Variable(self.detach().clone(), name=getattr(self, "_tf_name", None))


class IndexedSlices:
    def __init__(self, values, indices, dense_shape=None):
        self.values, self.indices, self.dense_shape = values, indices, dense_shape


class GradientTape:
    """Records embedding lookups so that `gradient` can return IndexedSlices."""

    def __init__(self, persistent=False, watch_accessed_variables=True):
        self.lookups = []     # (variable, flat int64 indices, gathered leaf rows)

    def __enter__(self):
        _ACTIVE_TAPES.append(self)
        return self

    def __exit__(self, *exc):
        _ACTIVE_TAPES.remove(self)
        return False

    def gradient(self, target, sources):
        sources = list(sources)
        leaves = [rows for (_, _, rows) in self.lookups] + sources
        if not isinstance(target, _torch.Tensor) or not target.requires_grad:
            return [None for _ in sources]
        grads = _torch.autograd.grad(target, leaves, allow_unused=True)
        nl = len(self.lookups)
        out = []
        for si, var in enumerate(sources):
            dense = grads[nl + si]
            vals, idxs = [], []
            for (v, idx, rows), g in zip(self.lookups, grads[:nl]):
                if v is var:
                    g = _torch.zeros_like(rows) if g is None else g
                    vals.append(g.reshape((-1,) + tuple(var.shape[1:])))
                    idxs.append(idx.reshape(-1))
            if dense is None and not vals:
                out.append(None)
            elif dense is None:
                out.append(IndexedSlices(_torch.cat(vals).as_subclass(Tensor), _torch.cat(idxs), tuple(var.shape)))
            else:
                d = dense.clone().as_subclass(_torch.Tensor)
                for val, idx in zip(vals, idxs):      # IndexedSlices + dense -> densified sum
                    d.index_add_(0, idx, val.as_subclass(_torch.Tensor))
                out.append(d.as_subclass(Tensor))
        return out


# ---- elementwise / shape ops ------------------------------------------------------
class _ClipByValue(_torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, lo, hi):
        ctx.save_for_backward((x >= lo) & (x <= hi))
        return x.clamp(lo, hi)

    @staticmethod
    def backward(ctx, g):
        (mask,) = ctx.saved_tensors
        return g * mask.to(g.dtype), None, None


def clip_by_value(t, clip_value_min, clip_value_max):
    return _ClipByValue.apply(_t(t), float(clip_value_min), float(clip_value_max)).as_subclass(Tensor)


def constant(value, dtype=None):
    return _t(_np.asarray(value), dtype)


def cast(x, dtype):
    return _t(x).to(_dt(dtype)).as_subclass(Tensor)


def range(*args, dtype=None):  # noqa: A001
    return _torch.arange(*args, dtype=_dt(dtype) or _torch.int32).as_subclass(Tensor)


def expand_dims(input, axis):  # noqa: A002
    return _t(input).unsqueeze(axis)


def squeeze(input, axis=None):  # noqa: A002
    x = _t(input)
    return x.squeeze() if axis is None else x.squeeze(axis)


def _axis(axis):
    return tuple(axis) if isinstance(axis, (list, tuple)) else axis


def reduce_sum(input_tensor, axis=None, keepdims=False):
    x = _t(input_tensor)
    return x.sum() if axis is None else x.sum(dim=_axis(axis), keepdim=keepdims)


def reduce_mean(input_tensor, axis=None, keepdims=False):
    x = _t(input_tensor)
    return x.mean() if axis is None else x.mean(dim=_axis(axis), keepdim=keepdims)


def reduce_max(input_tensor, axis=None, keepdims=False):
    x = _t(input_tensor)
    # torch.amax splits the gradient equally among ties, as tf.reduce_max does
    return x.amax() if axis is None else x.amax(dim=_axis(axis), keepdim=keepdims)


def pow(x, y):  # noqa: A001
    return _torch.pow(_t(x), y)


def abs(x):  # noqa: A001
    return _torch.abs(_t(x))


def exp(x):
    return _torch.exp(_t(x))


def multiply(x, y):
    return _t(x) * _t(y)


def matmul(a, b, transpose_a=False, transpose_b=False):
    a, b = _t(a), _t(b)
    if transpose_a:
        a = a.transpose(-1, -2)
    if transpose_b:
        b = b.transpose(-1, -2)
    return _torch.matmul(a, b)


def transpose(a, perm=None):
    a = _t(a)
    return a.permute(*perm) if perm is not None else a.permute(*reversed(list(_np.arange(a.dim()))))


def reshape(tensor, shape):
    return _t(tensor).reshape(tuple(shape))


def stack(values, axis=0):
    return _torch.stack([_t(v) for v in values], dim=axis)


def concat(values, axis):
    return _torch.cat([_t(v) for v in values], dim=axis)


def repeat(input, repeats, axis=None):  # noqa: A002
    x = _t(input)
    if axis is None:
        x = x.reshape(-1)
        axis = 0
    return _torch.repeat_interleave(x, int(repeats), dim=axis)


def eye(num_rows, num_columns=None, batch_shape=None, dtype=float32):
    e = _torch.eye(num_rows, num_columns if num_columns is not None else num_rows, dtype=_dt(dtype))
    if batch_shape:
        e = e.repeat(*batch_shape, 1, 1)
    return e.as_subclass(Tensor)


def complex(real, imag):  # noqa: A001
    return _torch.complex(_t(real), _t(imag))


def norm(tensor, ord="euclidean", axis=None, keepdims=False):
    x = _t(tensor)
    assert ord in ("euclidean", 2)
    s = (x * x).sum() if axis is None else (x * x).sum(dim=_axis(axis), keepdim=keepdims)
    return _torch.sqrt(s)


def stop_gradient(x):
    return _t(x).detach()


def gather_nd(params, indices):
    p, i = _t(params), _t(indices).long()
    assert i.shape[-1] == 1 and p.dim() == 1
    return p[i[..., 0]]


def tensor_scatter_nd_update(tensor, indices, updates):
    out = _t(tensor).clone()
    i = _t(indices).long()
    if i.numel() == 0:
        return out
    u = _torch.as_tensor(_np.asarray(updates, dtype=_np.float32))
    out.as_subclass(_torch.Tensor)[i[:, 0]] = u
    return out


def clip_by_norm(t, clip_norm, axes=None):
    """tf.clip_by_norm: t * clip / max(||t||_2, clip); IndexedSlices: norm over values."""
    if isinstance(t, IndexedSlices):
        return IndexedSlices(clip_by_norm(t.values, clip_norm), t.indices, t.dense_shape)
    x = _t(t)
    l2 = _torch.sqrt((x * x).sum())
    return x * clip_norm / _torch.maximum(l2, _torch.tensor(float(clip_norm)))


def _embedding_lookup(params, ids):
    idx = _t(ids).long().as_subclass(_torch.Tensor)
    if isinstance(params, Variable) and _ACTIVE_TAPES:
        rows = params.detach().as_subclass(_torch.Tensor)[idx].clone().requires_grad_(True)
        _ACTIVE_TAPES[-1].lookups.append((params, idx, rows))
        return rows.as_subclass(Tensor)
    return _t(params)[idx]


nn = _types.SimpleNamespace(
    embedding_lookup=_embedding_lookup,
    softmax=lambda logits, axis=-1: _torch.softmax(_t(logits), dim=axis),
)
math = _types.SimpleNamespace(
    sin=lambda x: _torch.sin(_t(x)), cos=lambda x: _torch.cos(_t(x)),
    log=lambda x: _torch.log(_t(x)), exp=exp,
    log_sigmoid=lambda x: _torch.nn.functional.logsigmoid(_t(x)),
)


# ---- random / initialisers --------------------------------------------------------
_GLOBAL_RNG = _np.random.default_rng(0)


def _set_seed(seed):
    global _GLOBAL_RNG
    _GLOBAL_RNG = _np.random.default_rng(seed)


def _random_uniform(shape, minval=0, maxval=None, dtype=float32, seed=None):
    dt = _dt(dtype)
    shape = tuple(int(s) for s in shape)
    if dt in (_torch.int32, _torch.int64):
        a = _GLOBAL_RNG.integers(int(minval), int(maxval), size=shape)
        return _torch.as_tensor(a).to(dt).as_subclass(Tensor)
    a = _GLOBAL_RNG.uniform(float(minval), float(maxval if maxval is not None else 1.0), size=shape)
    return _torch.as_tensor(a.astype(_np.float32)).as_subclass(Tensor)


random = _types.SimpleNamespace(uniform=_random_uniform, set_seed=_set_seed)


class _RandomUniform:
    """tf.initializers.RandomUniform(minval, maxval, seed): with a seed the values depend
    only on (seed, shape) — so equal shapes from one initialiser object are IDENTICAL
    (SURVEY.md C.3).  The stream itself is torch's, not TF's."""

    def __init__(self, minval=-0.05, maxval=0.05, seed=None):
        self.minval, self.maxval, self.seed = float(minval), float(maxval), seed

    def __call__(self, shape, dtype=None):
        g = _torch.Generator()
        if self.seed is not None:
            g.manual_seed(int(self.seed))
        else:
            g.seed()
        u = _torch.rand(tuple(int(s) for s in shape), generator=g, dtype=_torch.float32)
        return (u * (self.maxval - self.minval) + self.minval).as_subclass(Tensor)


initializers = _types.SimpleNamespace(RandomUniform=_RandomUniform)


# ---- Keras Adam (OptimizerV2) ---------------------------------
class _Adam:
    def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, amsgrad=False, name="Adam"):
        assert not amsgrad
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self.iterations = 0
        self._slots = {}

    def _mv(self, var):
        k = id(var)
        if k not in self._slots:
            z = _torch.zeros_like(var.detach().as_subclass(_torch.Tensor))
            self._slots[k] = (z, z.clone())
        return self._slots[k]

    def apply_gradients(self, grads_and_vars):
        self.iterations += 1
        f32 = _np.float32
        b1, b2 = f32(self.beta_1), f32(self.beta_2)
        b1p, b2p = b1 ** f32(self.iterations), b2 ** f32(self.iterations)
        lr_t = float(f32(self.learning_rate) * _np.sqrt(f32(1) - b2p) / (f32(1) - b1p))
        with _torch.no_grad():
            for g, var in grads_and_vars:
                if g is None:
                    continue
                m, v = self._mv(var)
                w = var.as_subclass(_torch.Tensor)
                if isinstance(g, IndexedSlices):
                    idx = g.indices
                    vals = g.values.detach().as_subclass(_torch.Tensor)
                    uniq, inv = _torch.unique(idx, return_inverse=True)
                    summed = _torch.zeros((len(uniq),) + tuple(vals.shape[1:]), dtype=vals.dtype)
                    summed.index_add_(0, inv, vals)
                    m.mul_(self.beta_1)
                    m[uniq] += (1 - self.beta_1) * summed
                    v.mul_(self.beta_2)
                    v[uniq] += (1 - self.beta_2) * summed * summed
                else:
                    gd = g.detach().as_subclass(_torch.Tensor)
                    m.mul_(self.beta_1).add_(gd, alpha=1 - self.beta_1)
                    v.mul_(self.beta_2).add_(gd * gd, alpha=1 - self.beta_2)
                w.sub_(lr_t * m / (_torch.sqrt(v) + self.epsilon))


class _SGD:
    """tf.optimizers.SGD, TF 2.5 gradient_descent.py: dense `var -= lr * g`; sparse (momentum = 0)
    `_resource_apply_sparse_duplicate_indices` = resource_scatter_add(var, indices, -lr * g): every
    IndexedSlices occurrence is applied, no de-duplication needed."""

    def __init__(self, learning_rate=0.01, momentum=0.0, nesterov=False, name="SGD"):
        assert momentum == 0.0 and not nesterov, "shim: plain SGD only"
        self.learning_rate, self.iterations = learning_rate, 0

    def apply_gradients(self, grads_and_vars):
        self.iterations += 1
        lr = float(_np.float32(self.learning_rate))
        with _torch.no_grad():
            for g, var in grads_and_vars:
                if g is None:
                    continue
                w = var.as_subclass(_torch.Tensor)
                if isinstance(g, IndexedSlices):
                    vals = g.values.detach().as_subclass(_torch.Tensor)
                    w.index_add_(0, g.indices, vals * (-lr))
                else:
                    w.sub_(g.detach().as_subclass(_torch.Tensor) * lr)


class _Adagrad:
    """tf.optimizers.Adagrad, TF 2.5 adagrad.py: slot `accumulator` = initial_accumulator_value; dense
    ResourceApplyAdagradV2 `acc += g*g; var -= lr * g / (sqrt(acc) + eps)`; sparse: OptimizerV2 first
    sums duplicate indices (_deduplicate_indexed_slices), then ResourceSparseApplyAdagradV2 does the same
    on the listed rows only (other rows are untouched)."""

    def __init__(self, learning_rate=0.001, initial_accumulator_value=0.1, epsilon=1e-7, name="Adagrad"):
        self.learning_rate, self.initial_accumulator_value, self.epsilon = learning_rate, initial_accumulator_value, epsilon
        self.iterations, self._acc = 0, {}

    def _slot(self, var):
        k = id(var)
        if k not in self._acc:
            self._acc[k] = _torch.full_like(var.detach().as_subclass(_torch.Tensor), self.initial_accumulator_value)
        return self._acc[k]

    def apply_gradients(self, grads_and_vars):
        self.iterations += 1
        lr, eps = float(_np.float32(self.learning_rate)), float(_np.float32(self.epsilon))
        with _torch.no_grad():
            for g, var in grads_and_vars:
                if g is None:
                    continue
                acc = self._slot(var)
                w = var.as_subclass(_torch.Tensor)
                if isinstance(g, IndexedSlices):
                    vals = g.values.detach().as_subclass(_torch.Tensor)
                    uniq, inv = _torch.unique(g.indices, return_inverse=True)
                    summed = _torch.zeros((len(uniq),) + tuple(vals.shape[1:]), dtype=vals.dtype)
                    summed.index_add_(0, inv, vals)
                    a = acc[uniq] + summed * summed
                    acc[uniq] = a
                    w[uniq] = w[uniq] - lr * summed / (_torch.sqrt(a) + eps)
                else:
                    gd = g.detach().as_subclass(_torch.Tensor)
                    acc.add_(gd * gd)
                    w.sub_(lr * gd / (_torch.sqrt(acc) + eps))


optimizers = _types.SimpleNamespace(Adam=_Adam, SGD=_SGD, Adagrad=_Adagrad)
keras = _types.SimpleNamespace(optimizers=optimizers)


# ---- tf.data: only from_tensor_slices(...).shuffle().repeat().batch().prefetch() -----
class _Dataset:
    def __init__(self, data):
        self.data = _np.asarray(data)
        self._shuffle = None
        self._batch = None

    @staticmethod
    def from_tensor_slices(data):
        return _Dataset(data)

    def shuffle(self, buffer_size, seed=None, reshuffle_each_iteration=True):
        assert buffer_size >= len(self.data), "shim supports full-buffer shuffles only"
        self._shuffle = _np.random.default_rng(seed)
        return self

    def repeat(self, count=None):
        return self

    def batch(self, batch_size):
        self._batch = int(batch_size)
        return self

    def prefetch(self, n):
        return self

    def __iter__(self):
        def rows():
            while True:    # repeat() precedes batch(): batches straddle passes (App. B.5)
                order = self._shuffle.permutation(len(self.data)) if self._shuffle is not None \
                    else _np.arange(len(self.data))
                for i in order:
                    yield self.data[i]
        it = rows()
        while True:
            yield _t(_np.stack([next(it) for _ in _builtin_range(self._batch)]))


import builtins as _builtins  # noqa: E402
_builtin_range = _builtins.range
data = _types.SimpleNamespace(Dataset=_Dataset, experimental=_types.SimpleNamespace(AUTOTUNE=-1))


# ---- inert stubs: logging / checkpoint / device config ------------------------------
class _Writer:
    def as_default(self):
        import contextlib
        return contextlib.nullcontext()


summary = _types.SimpleNamespace(create_file_writer=lambda *a, **k: _Writer(),
                                 scalar=lambda *a, **k: None, histogram=lambda *a, **k: None)


class _Checkpoint:
    def __init__(self, **kw):
        pass

    def restore(self, *a, **k):
        return self


class _CheckpointManager:
    latest_checkpoint = None

    def __init__(self, checkpoint=None, directory=None, max_to_keep=None):
        pass

    def save(self, *a, **k):
        return None


train = _types.SimpleNamespace(Checkpoint=_Checkpoint, CheckpointManager=_CheckpointManager)
config = _types.SimpleNamespace(
    list_physical_devices=lambda kind=None: [],
    experimental=_types.SimpleNamespace(set_memory_growth=lambda *a, **k: None))
