"""tensorflow.keras.layers of the NumPy stand-in (TEST INFRASTRUCTURE ONLY)."""

from __future__ import annotations

import numpy as np

_ALL_LAYERS = []          # creation order


class KerasTensor:
    """A symbolic tensor of the functional API: a thunk over the Input placeholders."""

    def __init__(self, fn, name=None, shape=None, parents=(), layer=None):
        self.fn, self.name, self._shape, self.parents, self.layer = fn, name, shape, tuple(parents), layer

    def __getitem__(self, idx):
        return KerasTensor(lambda env, cache: resolve(self, env, cache)[idx], parents=[self])


def is_symbolic(x):
    if isinstance(x, KerasTensor):
        return True
    if isinstance(x, (list, tuple)):
        return any(is_symbolic(v) for v in x)
    return False


def resolve(x, env, cache):
    if isinstance(x, KerasTensor):
        if id(x) in env:
            return env[id(x)]
        if id(x) not in cache:
            cache[id(x)] = x.fn(env, cache)
        return cache[id(x)]
    if isinstance(x, (list, tuple)):
        return type(x)(resolve(v, env, cache) for v in x)
    return x


def collect_layers(outputs):
    seen, found = set(), []

    def walk(t):
        if isinstance(t, (list, tuple)):
            for v in t:
                walk(v)
            return
        if not isinstance(t, KerasTensor) or id(t) in seen:
            return
        seen.add(id(t))
        if t.layer is not None and t.layer not in found:
            found.append(t.layer)
        for p in t.parents:
            walk(p)
    walk(outputs)
    return sorted(found, key=lambda l: l._creation_index)


def symbolic_op(fn):
    """Wrap a tf function so that symbolic arguments defer it."""
    def op(*args, **kwargs):
        if is_symbolic(args) or is_symbolic(list(kwargs.values())):
            parents = [a for a in _flatten(args) if isinstance(a, KerasTensor)]
            return KerasTensor(lambda env, cache: fn(*resolve(list(args), env, cache), **{k: resolve(v, env, cache) for k, v in kwargs.items()}),
                               parents=parents)
        return fn(*args, **kwargs)
    return op


def _flatten(x):
    for v in x:
        if isinstance(v, (list, tuple)):
            yield from _flatten(v)
        else:
            yield v


class Layer:
    def __init__(self, name=None, **kwargs):
        self.name = name
        self._creation_index = len(_ALL_LAYERS)
        _ALL_LAYERS.append(self)

    @property
    def trainable_variables(self):
        return []

    def call(self, inputs):
        raise NotImplementedError

    def __call__(self, inputs, *args, **kwargs):
        if is_symbolic(inputs):
            parents = [a for a in _flatten([inputs]) if isinstance(a, KerasTensor)]
            return KerasTensor(lambda env, cache: self.call(resolve(inputs, env, cache)), parents=parents, layer=self)
        return self.call(inputs)


class Dense(Layer):
    """y = activation(x @ kernel + bias); kernel [in, units] created at the first call (Glorot uniform) unless set."""

    def __init__(self, units, activation=None, name=None, **kwargs):
        super().__init__(name=name)
        self.units, self.activation = int(units), activation
        self.kernel = self.bias = None

    def set_weights(self, weights):
        self.kernel, self.bias = np.asarray(weights[0], np.float32), np.asarray(weights[1], np.float32)

    def get_weights(self):
        return [self.kernel, self.bias]

    @property
    def trainable_variables(self):
        return [self.kernel, self.bias]

    def call(self, inputs):
        import tensorflow as tf
        x = np.asarray(inputs.numpy(), np.float32)
        if self.kernel is None:
            lim = np.sqrt(6.0 / (x.shape[-1] + self.units))
            self.kernel = np.random.default_rng(self._creation_index).uniform(-lim, lim, (x.shape[-1], self.units)).astype(np.float32)
            self.bias = np.zeros(self.units, np.float32)
        assert self.kernel.shape[0] == x.shape[-1], (self.kernel.shape, x.shape)
        # fp32 inputs and weights, products accumulated in float64 and rounded once: the reference arithmetic the parity
        # bars are written against is "fp32 Dense", whose summation order inside the GEMM is TensorFlow's own business
        y = (x.astype(np.float64) @ self.kernel.astype(np.float64) + self.bias.astype(np.float64)).astype(np.float32)
        if self.activation == "relu":
            y = np.maximum(y, np.float32(0))
        elif self.activation is not None:
            raise NotImplementedError(self.activation)
        return tf.Tensor(y)


class LeakyReLU(Layer):
    def __init__(self, alpha=0.3, **kwargs):
        super().__init__()
        self.alpha = np.float32(alpha)

    def call(self, inputs):
        import tensorflow as tf
        a = inputs.numpy()
        return tf.Tensor(np.where(a > 0, a, a * self.alpha))
