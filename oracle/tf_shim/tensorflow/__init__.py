"""TEST INFRASTRUCTURE ONLY (see oracle/README.md): a NumPy stand-in for the part of the TensorFlow 2 API that the
reference's network/*.py and util/interpolate.py call, so that the reference's OWN Python source — chunking, culling
and scatter, stratified and hierarchical sampling, the compositing expressions, the instanced and mip renderers, the
functional Keras model graphs — can be imported from /root/reference and executed here without TensorFlow.

This is OUR code, not TensorFlow: every op is the documented semantics of the TF op of the same name written with
NumPy in float32 (eager only; no graphs, no gradients, no devices).  Ops whose exact arithmetic is part of the parity
contract follow the same restatement as oracle/nerftex_oracle.py — `linspace` (start + i * delta with delta rounded
once, exact end point), `cumsum` / `cumprod` / `reduce_sum` (left to right in fp32), `searchsorted(side='right')`
(upper bound) — so what running the reference over this pins is the reference's LOGIC (which tensors, which slices,
which order of operations), not TensorFlow's kernels.

`random.uniform` / `random.normal` draw from a seedable generator and append every draw to `random.log`, so a test
can hand the very same numbers to the code it compares against."""

from __future__ import annotations

import builtins as _builtins
import contextlib

import numpy as np

from . import _autodiff as _ad
from . import keras  # noqa: F401  (tensorflow.keras)
from ._autodiff import GradientTape  # noqa: F401

__version__ = "2.4.1-numpy-stand-in"

float32 = np.float32
float64 = np.float64
int32 = np.int32
int64 = np.int64
bool = np.bool_  # noqa: A001


class dtypes:
    float32 = np.float32
    float64 = np.float64
    int32 = np.int32
    int64 = np.int64


class Shape(list):
    """TensorShape: a list that stays a Shape under slicing and `+ list`."""

    def __getitem__(self, i):
        r = list.__getitem__(self, i)
        return Shape(r) if isinstance(i, slice) else r

    def __add__(self, other):
        return Shape(list(self) + list(other))

    def __radd__(self, other):
        return Shape(list(other) + list(self))

    def as_list(self):
        return list(self)


def _np(x, dtype=None):
    """ndarray of a Tensor / ndarray / Python value; Python floats become float32 like tf.convert_to_tensor makes them."""
    if isinstance(x, Tensor):
        a = x._a
    elif isinstance(x, np.ndarray):
        a = x
    elif isinstance(x, (list, tuple)):
        a = np.asarray([_np(v) for v in x]) if any(isinstance(v, Tensor) for v in x) else np.asarray(x)
        if a.dtype == np.float64:
            a = a.astype(np.float32)
        if a.dtype == np.int64 and dtype is None:
            a = a.astype(np.int32)
    elif isinstance(x, float):
        a = np.float32(x)
    else:
        a = np.asarray(x)
    return a if dtype is None else np.asarray(a, dtype)


def _pair(a, b):
    """Operands of a binary op: a Python scalar takes the dtype of the tensor it meets."""
    ta, tb = isinstance(a, (Tensor, np.ndarray)), isinstance(b, (Tensor, np.ndarray))
    x, y = _np(a), _np(b)
    if ta and not tb:
        y = np.asarray(y, x.dtype) if np.ndim(y) == 0 else y
    elif tb and not ta:
        x = np.asarray(x, y.dtype) if np.ndim(x) == 0 else x
    return x, y


class Tensor:
    __array_priority__ = 100

    _node = None                                   # set by _autodiff.attach while a GradientTape records

    def __init__(self, a):
        self._a = np.asarray(a)

    # --- introspection
    @property
    def shape(self):
        return Shape(self._a.shape)

    @property
    def dtype(self):
        return self._a.dtype.type

    @property
    def ndim(self):
        return self._a.ndim

    def numpy(self):
        return self._a

    def __array__(self, dtype=None, copy=None):
        return self._a if dtype is None else self._a.astype(dtype)

    def __len__(self):
        return self._a.shape[0]

    def __iter__(self):
        for i in _builtins.range(self._a.shape[0]):
            yield Tensor(self._a[i])

    def __repr__(self):
        return f"<stand-in tf.Tensor shape={tuple(self._a.shape)} dtype={self._a.dtype}>"

    def __float__(self):
        return float(self._a)

    def __int__(self):
        return int(self._a)

    def __index__(self):
        return int(self._a)

    def __bool__(self):
        return builtins_bool(self._a)

    __hash__ = object.__hash__

    # --- indexing
    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        idx = tuple(int(i._a) if isinstance(i, Tensor) and i._a.ndim == 0 else (i._a if isinstance(i, Tensor) else i) for i in idx)
        idx = tuple(slice(*[None if v is None else int(_np(v)) for v in (i.start, i.stop, i.step)]) if isinstance(i, slice) else i
                    for i in idx)
        out = Tensor(self._a[idx])
        if _ad.recording(self):
            shp, dt = self._a.shape, self._a.dtype

            def vjp(g, idx=idx):
                z = np.zeros(shp, dt)
                np.add.at(z, idx, g)
                return (z,)
            _ad.attach(out, (self,), vjp)
        return out

    # --- arithmetic (float32 stays float32; Python scalars adopt the tensor's dtype)
    def _bin(self, other, fn, swap=False):
        a_, b_ = (other, self) if swap else (self, other)
        x, y = _pair(a_, b_)
        with np.errstate(all="ignore"):
            out = Tensor(fn(x, y))
        if _ad.recording(a_, b_):
            xs, ys = np.shape(x), np.shape(y)
            if fn is np.add:
                vjp = lambda g: (_ad.unbroadcast(g, xs), _ad.unbroadcast(g, ys))
            elif fn is np.subtract:
                vjp = lambda g: (_ad.unbroadcast(g, xs), _ad.unbroadcast(-g, ys))
            elif fn is np.multiply:
                vjp = lambda g: (_ad.unbroadcast(g * y, xs), _ad.unbroadcast(g * x, ys))
            elif fn is np.divide:
                vjp = lambda g: (_ad.unbroadcast(g / y, xs), _ad.unbroadcast(-g * x / (y * y), ys))
            elif fn is np.power and not _ad.tracked(b_):
                vjp = lambda g: (_ad.unbroadcast(g * y * np.power(x, y - 1), xs), None)
            else:
                _ad.DROPPED.add(getattr(fn, "__name__", str(fn)))
                return out
            _ad.attach(out, (a_ if isinstance(a_, Tensor) else None, b_ if isinstance(b_, Tensor) else None), vjp)
        return out

    def __add__(self, o): return self._bin(o, np.add)
    def __radd__(self, o): return self._bin(o, np.add, True)
    def __sub__(self, o): return self._bin(o, np.subtract)
    def __rsub__(self, o): return self._bin(o, np.subtract, True)
    def __mul__(self, o): return self._bin(o, np.multiply)
    def __rmul__(self, o): return self._bin(o, np.multiply, True)
    def __truediv__(self, o): return self._bin(o, np.divide)
    def __rtruediv__(self, o): return self._bin(o, np.divide, True)
    def __pow__(self, o): return self._bin(o, np.power)
    def __rpow__(self, o): return self._bin(o, np.power, True)
    def __mod__(self, o): return self._bin(o, np.mod)
    def __floordiv__(self, o): return self._bin(o, np.floor_divide)
    def __rfloordiv__(self, o): return self._bin(o, np.floor_divide, True)
    def __neg__(self):
        out = Tensor(-self._a)
        return _ad.attach(out, (self,), lambda g: (-g,)) if _ad.recording(self) else out
    def __abs__(self): return Tensor(np.abs(self._a))
    def __lt__(self, o): return self._bin(o, np.less)
    def __le__(self, o): return self._bin(o, np.less_equal)
    def __gt__(self, o): return self._bin(o, np.greater)
    def __ge__(self, o): return self._bin(o, np.greater_equal)
    def __eq__(self, o): return self._bin(o, np.equal)
    def __ne__(self, o): return self._bin(o, np.not_equal)
    def __and__(self, o): return self._bin(o, np.logical_and)
    def __or__(self, o): return self._bin(o, np.logical_or)
    def __invert__(self): return Tensor(np.logical_not(self._a))


builtins_bool = _builtins.bool


def _t(a):
    return Tensor(a)


# ------------------------------------------------------------------------------------------------ creation / shape
def constant(value, dtype=None, shape=None):
    a = _np(value, dtype)
    return _t(a if shape is None else np.broadcast_to(a, shape).copy())


convert_to_tensor = constant


def cast(x, dtype):
    out = _t(_np(x).astype(dtype))
    if _ad.recording(x) and np.issubdtype(dtype, np.floating):
        _ad.attach(out, (x,), lambda g: (g,))
    return out


def zeros(shape, dtype=np.float32):
    return _t(np.zeros([int(s) for s in (shape if isinstance(shape, (list, tuple)) else [shape])], dtype))


def ones(shape, dtype=np.float32):
    return _t(np.ones([int(s) for s in (shape if isinstance(shape, (list, tuple)) else [shape])], dtype))


def zeros_like(x):
    return _t(np.zeros_like(_np(x)))


def ones_like(x):
    return _t(np.ones_like(_np(x)))


def range(start, limit=None, delta=1, dtype=None):  # noqa: A001
    if limit is None:
        start, limit = 0, start
    a = np.arange(_np(start), _np(limit), _np(delta))
    if dtype is not None:
        a = a.astype(dtype)
    elif a.dtype == np.int64:
        a = a.astype(np.int32)
    elif a.dtype == np.float64:
        a = a.astype(np.float32)
    return _t(a)


def linspace(start, stop, num):
    """tf.linspace: start + i * delta, delta = (stop - start) / (num - 1) rounded once to fp32, the end point exact."""
    start, stop, num = np.float32(start), np.float32(stop), int(num)
    if num == 1:
        return _t(np.asarray([start], np.float32))
    delta = np.float32((stop - start) / np.float32(num - 1))
    a = (start + (np.arange(num, dtype=np.float32) * delta).astype(np.float32)).astype(np.float32)
    a[-1] = stop
    return _t(a)


def shape(x):
    return _t(np.asarray(_np(x).shape, np.int32))


def rank(x):
    return _t(np.asarray(_np(x).ndim, np.int32))


def reshape(x, shp):
    shp = [shp] if isinstance(shp, (int, np.integer)) else [int(_np(s)) for s in shp]
    out = _t(_np(x).reshape(shp))
    if _ad.recording(x):
        old = _np(x).shape
        _ad.attach(out, (x,), lambda g: (g.reshape(old),))
    return out


def expand_dims(x, axis):
    out = _t(np.expand_dims(_np(x), axis))
    if _ad.recording(x):
        old = _np(x).shape
        _ad.attach(out, (x,), lambda g: (g.reshape(old),))
    return out


def concat(values, axis):
    arrs = [_np(v) for v in values]
    out = _t(np.concatenate(arrs, axis))
    if _ad.recording(*values):
        cuts = np.cumsum([a.shape[axis] for a in arrs])[:-1]
        _ad.attach(out, tuple(v if isinstance(v, Tensor) else None for v in values), lambda g: tuple(np.split(g, cuts, axis)))
    return out


def stack(values, axis=0):
    _ad.dropped("stack", *values)
    return _t(np.stack([_np(v) for v in values], axis))


def repeat(x, repeats, axis=None):
    _ad.dropped("repeat", x)
    return _t(np.repeat(_np(x), int(_np(repeats)), axis))


def broadcast_to(x, shp):
    out = _t(np.broadcast_to(_np(x), [int(_np(s)) for s in shp]).copy())
    if _ad.recording(x):
        old = _np(x).shape
        _ad.attach(out, (x,), lambda g: (_ad.unbroadcast(g, old),))
    return out


def stop_gradient(x):
    return _t(_np(x)) if isinstance(x, Tensor) else x


def sort(x, axis=-1, direction="ASCENDING"):
    _ad.dropped("sort", x)
    a = np.sort(_np(x), axis=axis, kind="stable")
    return _t(a if direction == "ASCENDING" else np.flip(a, axis))


# ------------------------------------------------------------------------------------------------ gather / scatter
def where(condition, x=None, y=None):
    c = _np(condition)
    if x is None and y is None:
        return _t(np.argwhere(c).astype(np.int64))                 # [n, rank] coordinates, row-major order
    _ad.dropped("where", x, y)
    a, b = _pair(x, y)
    return _t(np.where(c, a, b))


def gather_nd(params, indices):
    p, i = _np(params), _np(indices)
    key = tuple(i[..., k] for k in np.arange(i.shape[-1]))
    out = _t(p[key])
    if _ad.recording(params):
        def vjp(g):
            z = np.zeros(p.shape, p.dtype)
            np.add.at(z, key, g)
            return (z,)
        _ad.attach(out, (params,), vjp)
    return out


def scatter_nd(indices, updates, shape):  # noqa: A002
    i, u = _np(indices), _np(updates)
    out = np.zeros([int(_np(s)) for s in shape], u.dtype)
    key = tuple(i[..., k] for k in np.arange(i.shape[-1]))
    np.add.at(out, key, u)                                             # duplicates add up, like tf.scatter_nd
    out = _t(out)
    if _ad.recording(updates):
        _ad.attach(out, (updates,), lambda g: (g[key],))
    return out


def gather(params, indices, axis=None, batch_dims=0):
    _ad.dropped("gather", params)
    p, i = _np(params), _np(indices)
    if batch_dims == 0:
        return _t(np.take(p, i, axis=0 if axis is None else axis))
    axis = p.ndim - 1 if axis in (-1, None) else axis
    assert axis == batch_dims == p.ndim - 1, "stand-in: gather along the last axis with all leading axes as batch"
    flat_p = p.reshape(-1, p.shape[-1])
    flat_i = i.reshape(flat_p.shape[0], -1)
    out = np.take_along_axis(flat_p, flat_i, axis=1)
    return _t(out.reshape(i.shape))


def searchsorted(sorted_sequence, values, side="left"):
    _ad.dropped("searchsorted", sorted_sequence, values)
    s, v = _np(sorted_sequence), _np(values)
    fs, fv = s.reshape(-1, s.shape[-1]), v.reshape(-1, v.shape[-1])
    out = np.stack([np.searchsorted(fs[r], fv[r], side=side) for r in np.arange(fs.shape[0])]) if fs.shape[0] else np.zeros(fv.shape, np.int64)
    return _t(out.reshape(v.shape).astype(np.int32))


# ------------------------------------------------------------------------------------------------ reductions / scans
def _seq_reduce(a, axis, fn):
    """Left-to-right fp32 reduction along `axis` (the order the parity contract defines)."""
    a = np.moveaxis(a, axis, 0)
    acc = a[0].copy()
    for k in np.arange(1, a.shape[0]):
        acc = fn(acc, a[k]).astype(a.dtype)
    return acc


def reduce_sum(x, axis=None, keepdims=False):
    a = _np(x)
    if axis is None:
        out = _t(a.sum(dtype=a.dtype))
        return _ad.attach(out, (x,), lambda g: (np.broadcast_to(g, a.shape),)) if _ad.recording(x) else out
    axis = axis % a.ndim
    r = _seq_reduce(a, axis, np.add) if a.shape[axis] > 0 else np.zeros(np.delete(a.shape, axis), a.dtype)
    out = _t(np.expand_dims(r, axis) if keepdims else r)
    if _ad.recording(x):
        _ad.attach(out, (x,), lambda g: (np.broadcast_to(g if keepdims else np.expand_dims(g, axis), a.shape),))
    return out


def reduce_prod(x, axis=None, keepdims=False):
    return _t(np.prod(_np(x), axis=axis, keepdims=keepdims))


def reduce_mean(x, axis=None, keepdims=False):
    a = _np(x)
    out = _t(a.mean(axis=axis, keepdims=keepdims, dtype=a.dtype))
    if _ad.recording(x):
        assert axis is None, "stand-in: gradient of reduce_mean over all axes only"
        _ad.attach(out, (x,), lambda g: (np.broadcast_to(g / a.dtype.type(a.size), a.shape),))
    return out


def reduce_max(x, axis=None, keepdims=False):
    return _t(np.max(_np(x), axis=axis, keepdims=keepdims))


def reduce_min(x, axis=None, keepdims=False):
    return _t(np.min(_np(x), axis=axis, keepdims=keepdims))


def reduce_any(x, axis=None, keepdims=False):
    return _t(np.any(_np(x), axis=axis, keepdims=keepdims))


def cumsum(x, axis=0, exclusive=False, reverse=False):
    _ad.dropped("cumsum", x)
    a = _np(x)
    assert not reverse
    a = np.moveaxis(a, axis, 0)
    out = np.empty_like(a)
    acc = np.zeros_like(a[0])
    for k in np.arange(a.shape[0]):
        if exclusive:
            out[k] = acc
            acc = (acc + a[k]).astype(a.dtype)
        else:
            acc = (acc + a[k]).astype(a.dtype)
            out[k] = acc
    return _t(np.moveaxis(out, 0, axis))


def _cumprod(x, axis=0, exclusive=False, reverse=False):
    a = _np(x)
    assert not reverse
    a = np.moveaxis(a, axis, 0)
    out = np.empty_like(a)
    acc = np.ones_like(a[0])
    for k in np.arange(a.shape[0]):
        if exclusive:
            out[k] = acc
            acc = (acc * a[k]).astype(a.dtype)
        else:
            acc = (acc * a[k]).astype(a.dtype)
            out[k] = acc
    res = _t(np.moveaxis(out, 0, axis))
    if _ad.recording(x):
        assert exclusive, "stand-in: gradient of the exclusive cumprod only"

        def vjp(g):
            # y_k = prod_{j<k} x_j  =>  dL/dx_j = sum_{k>j} g_k y_k / x_j   (x = 1 - alpha + 1e-10 > 0 on this path)
            gy = np.moveaxis(g, axis, 0) * out
            suffix = np.cumsum(gy[::-1], axis=0)[::-1]
            excl = np.concatenate([suffix[1:], np.zeros_like(suffix[:1])], 0)
            return (np.moveaxis((excl / a).astype(a.dtype), 0, axis),)
        _ad.attach(res, (x,), vjp)
    return res


def tensordot(a, b, axes):
    return _t(np.tensordot(_np(a), _np(b), axes).astype(np.result_type(_np(a), _np(b))))


def foldl(fn, elems, initializer=None):
    it = iter(elems)
    acc = next(it) if initializer is None else initializer
    for e in it:
        acc = fn(acc, e)
    return acc


def map_fn(fn, elems, **kwargs):
    return stack([fn(e) for e in elems], 0)


# ------------------------------------------------------------------------------------------------ elementwise
def _unary(fn):
    def op(x, name=None):
        with np.errstate(all="ignore"):
            a = _np(x)
            return _t(fn(a).astype(a.dtype) if a.dtype.kind == "f" else fn(a))
    return op


def _unary64(fn):
    """Transcendentals: evaluated in float64 and rounded once, like oracle/nerftex_oracle.py does."""
    def op(x, name=None):
        with np.errstate(all="ignore"):
            a = _np(x)
            return _t(fn(a.astype(np.float64)).astype(a.dtype if a.dtype.kind == "f" else np.float32))
    return op


def exp(x, name=None):
    a = _np(x)
    with np.errstate(all="ignore"):
        y = np.exp(a.astype(np.float64)).astype(a.dtype if a.dtype.kind == "f" else np.float32)
    out = _t(y)
    return _ad.attach(out, (x,), lambda g: (g * y,)) if _ad.recording(x) else out


sin = _unary64(np.sin)
cos = _unary64(np.cos)
atan = _unary64(np.arctan)
def abs(x, name=None):  # noqa: A001
    a = _np(x)
    out = _t(np.abs(a))
    return _ad.attach(out, (x,), lambda g: (g * np.sign(a),)) if _ad.recording(x) else out


sqrt = _unary(np.sqrt)
floor = _unary(np.floor)


def _sigmoid(x):
    a = _np(x)
    with np.errstate(all="ignore"):
        y = (1.0 / (1.0 + np.exp(-a.astype(np.float64)))).astype(a.dtype)
    out = _t(y)
    return _ad.attach(out, (x,), lambda g: (g * y * (1 - y),)) if _ad.recording(x) else out


def maximum(x, y):
    _ad.dropped("maximum", x, y)
    return _t(np.maximum(*_pair(x, y)))


def minimum(x, y):
    _ad.dropped("minimum", x, y)
    return _t(np.minimum(*_pair(x, y)))


class math:
    sin, cos, exp, atan, abs, sqrt, floor = sin, cos, exp, atan, abs, sqrt, floor
    sigmoid = staticmethod(_sigmoid)
    cumprod = staticmethod(_cumprod)
    cumsum = staticmethod(cumsum)
    reduce_mean = staticmethod(reduce_mean)
    reduce_sum = staticmethod(reduce_sum)
    maximum, minimum = staticmethod(maximum), staticmethod(minimum)

    @staticmethod
    def squared_difference(x, y):
        a, b = _pair(x, y)
        d = a - b
        out = _t(d * d)
        if _ad.recording(x, y):
            _ad.attach(out, (x if isinstance(x, Tensor) else None, y if isinstance(y, Tensor) else None),
                       lambda g: (_ad.unbroadcast(2 * g * d, np.shape(a)), _ad.unbroadcast(-2 * g * d, np.shape(b))))
        return out

    @staticmethod
    def mod(x, y):
        return _t(np.mod(*_pair(x, y)))


class nn:
    @staticmethod
    def relu(x):
        a = _np(x)
        out = _t(np.maximum(a, a.dtype.type(0)))
        return _ad.attach(out, (x,), lambda g: (g * (a > 0),)) if _ad.recording(x) else out

    @staticmethod
    def elu(x):
        a = _np(x)
        with np.errstate(all="ignore"):
            e = np.expm1(a.astype(np.float64)).astype(a.dtype)
        out = _t(np.where(a > 0, a, e))
        return _ad.attach(out, (x,), lambda g: (g * np.where(a > 0, a.dtype.type(1), e + 1),)) if _ad.recording(x) else out

    @staticmethod
    def depthwise_conv2d(input, filter, strides, padding, **kwargs):  # noqa: A002
        """NHWC input, [kh, kw, C, 1] filter, 'SAME' padding as TensorFlow documents it: out = ceil(in / stride),
        pad_total = max((out - 1) * stride + k - in, 0), the smaller half in front.  Products accumulated in float64."""
        x, f = _np(input).astype(np.float64), _np(filter).astype(np.float64)
        assert padding == "SAME" and f.shape[3] == 1 and x.shape[3] == f.shape[2]
        n, H, W, C = x.shape
        kh, kw = f.shape[:2]
        sh, sw = int(strides[1]), int(strides[2])
        oh, ow = -(-H // sh), -(-W // sw)
        ph, pw = max((oh - 1) * sh + kh - H, 0), max((ow - 1) * sw + kw - W, 0)
        xp = np.zeros((n, H + ph, W + pw, C))
        xp[:, ph // 2:ph // 2 + H, pw // 2:pw // 2 + W] = x
        out = np.zeros((n, oh, ow, C))
        for i in np.arange(kh):
            for j in np.arange(kw):
                out += xp[:, i:i + (oh - 1) * sh + 1:sh, j:j + (ow - 1) * sw + 1:sw] * f[i, j, :, 0]
        return _t(out.astype(np.float32))


class linalg:
    @staticmethod
    def norm(x, axis=None, keepdims=False):
        _ad.dropped("norm", x)
        a = _np(x)
        sq = reduce_sum(_t(a * a), axis=axis, keepdims=keepdims)._a
        return _t(np.sqrt(sq).astype(a.dtype))

    @staticmethod
    def normalize(x, axis=None):
        a = _np(x)
        n = np.sqrt((a * a).sum(axis=axis, keepdims=True, dtype=a.dtype)).astype(a.dtype)
        return _t((a / n).astype(a.dtype)), _t(n)

    @staticmethod
    def cross(a, b):
        x, y = _pair(a, b)
        x, y = np.asarray(x, np.float32), np.asarray(y, np.float32)
        return _t(np.cross(x, y).astype(np.float32))


class image:
    @staticmethod
    def resize(images, size, method="bilinear", **kwargs):
        """'nearest' only, as TF 2 documents it (half-pixel centres): source = min(floor((i + 0.5) * in / out), in - 1)."""
        assert method == "nearest"
        a = _np(images)
        n, H, W, C = a.shape
        oh, ow = int(size[0]), int(size[1])
        ys = np.minimum(np.floor((np.arange(oh) + 0.5) * (H / oh)).astype(np.int64), H - 1)
        xs = np.minimum(np.floor((np.arange(ow) + 0.5) * (W / ow)).astype(np.int64), W - 1)
        return _t(a[:, ys][:, :, xs])


class bitwise:
    @staticmethod
    def right_shift(x, y):
        return _t(np.right_shift(_np(x), _np(y)))


class debugging:
    @staticmethod
    def check_numerics(x, message):
        if not np.all(np.isfinite(_np(x))):
            raise FloatingPointError(message)
        return x

    @staticmethod
    def Assert(condition, data=None):
        assert builtins_bool(np.all(_np(condition))), data


class _Random:
    def __init__(self):
        self.rng = np.random.default_rng(0)
        self.log = []

    def seed(self, s):
        self.rng = np.random.default_rng(s)
        self.log = []

    def set_seed(self, s):
        self.seed(s)

    def uniform(self, shape, minval=0, maxval=None, dtype=np.float32):  # noqa: A002
        shape = [int(_np(s)) for s in shape]
        if np.issubdtype(dtype, np.integer):
            a = self.rng.integers(minval, maxval, size=shape).astype(dtype)
        else:
            hi = 1.0 if maxval is None else maxval
            a = (self.rng.random(size=shape, dtype=np.float32) * np.float32(hi - minval) + np.float32(minval)).astype(np.float32)
        self.log.append(("uniform", a))
        return _t(a)

    def shuffle(self, value):
        a = _np(value)
        perm = self.rng.permutation(a.shape[0])
        self.log.append(("shuffle", perm))
        return _t(a[perm])

    def normal(self, shape, mean=0.0, stddev=1.0, dtype=np.float32):  # noqa: A002
        a = (self.rng.standard_normal(size=[int(_np(s)) for s in shape], dtype=np.float32) * np.float32(stddev) + np.float32(mean)).astype(np.float32)
        self.log.append(("normal", a))
        return _t(a)


random = _Random()


@contextlib.contextmanager
def name_scope(name):
    yield


# the functional Keras API writes graphs with these on symbolic tensors (network/model.py:28,44,85,93,101,108,115)
concat = keras.layers.symbolic_op(concat)


class Variable(Tensor):
    """A trainable leaf: always tracked while a tape records."""
    _node = _ad.LEAF

    def __init__(self, initial_value, dtype=None, trainable=True, name=None):
        super().__init__(np.array(_np(initial_value), dtype=dtype))
        self.name = name

    def assign(self, value):
        self._a = np.array(_np(value), dtype=self._a.dtype).reshape(self._a.shape)
        return self

    def assign_sub(self, value):
        self._a = (self._a - _np(value)).astype(self._a.dtype)
        return self

    def assign_add(self, value):
        self._a = (self._a + _np(value)).astype(self._a.dtype)
        return self


class data:
    class Dataset:
        pass


def function(fn=None, **kwargs):
    return fn if fn is not None else (lambda f: f)
