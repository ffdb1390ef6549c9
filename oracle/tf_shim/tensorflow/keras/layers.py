"""tensorflow.keras.layers of the NumPy stand-in (TEST INFRASTRUCTURE ONLY)."""

from __future__ import annotations

import numpy as np

_ALL_LAYERS = []          # creation order


class KerasTensor:
    """A symbolic tensor of the functional API: a thunk over the Input placeholders.  `dummy` is the op applied to one
    all-zero row (shape inference the way Keras builds layers when the graph is written: Dense creates its variables
    here, so that model.trainable_variables exists before the first real call, network/train.py:33-34)."""

    def __init__(self, fn, name=None, shape=None, parents=(), layer=None, dummy=None):
        self.fn, self.name, self._shape, self.parents, self.layer = fn, name, shape, tuple(parents), layer
        if dummy is None and fn is None:                              # an Input placeholder
            import tensorflow as tf
            last = 0 if (shape is None or shape[-1] is None) else int(shape[-1])
            dummy = tf.Tensor(np.zeros((1, last), np.float32))
        self.dummy = dummy

    def __getitem__(self, idx):
        return KerasTensor(lambda env, cache: resolve(self, env, cache)[idx], parents=[self], dummy=self.dummy[idx])


def dummies(x):
    if isinstance(x, KerasTensor):
        return x.dummy
    if isinstance(x, (list, tuple)):
        return type(x)(dummies(v) for v in x)
    return x


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
                               parents=parents, dummy=fn(*dummies(list(args)), **{k: dummies(v) for k, v in kwargs.items()}))
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
            return KerasTensor(lambda env, cache: self.call(resolve(inputs, env, cache)), parents=parents, layer=self,
                               dummy=self.call(dummies(inputs)))
        return self.call(inputs)


class Dense(Layer):
    """y = activation(x @ kernel + bias); kernel [in, units] created at the first call (Glorot uniform) unless set."""

    def __init__(self, units, activation=None, name=None, **kwargs):
        super().__init__(name=name)
        self.units, self.activation = int(units), activation
        self.kernel = self.bias = None

    def set_weights(self, weights):
        import tensorflow as tf
        if self.kernel is not None:                                   # in place: optimizers and checkpoints hold the objects
            self.kernel.assign(weights[0])
            self.bias.assign(weights[1])
            return
        self.kernel = tf.Variable(np.asarray(weights[0], np.float32), name=f"{self.name}/kernel")
        self.bias = tf.Variable(np.asarray(weights[1], np.float32), name=f"{self.name}/bias")

    def get_weights(self):
        return [self.kernel.numpy(), self.bias.numpy()]

    @property
    def trainable_variables(self):
        return [self.kernel, self.bias]

    def call(self, inputs):
        import tensorflow as tf
        from tensorflow import _autodiff as ad
        x = np.asarray(inputs.numpy(), np.float32)
        if self.kernel is None:
            lim = np.sqrt(6.0 / (x.shape[-1] + self.units))
            self.set_weights([np.random.default_rng(self._creation_index).uniform(-lim, lim, (x.shape[-1], self.units)),
                              np.zeros(self.units)])
        W, b = self.kernel.numpy(), self.bias.numpy()
        assert W.shape[0] == x.shape[-1], (W.shape, x.shape)
        # fp32 inputs and weights, products accumulated in float64 and rounded once: the reference arithmetic the parity
        # bars are written against is "fp32 Dense", whose summation order inside the GEMM is TensorFlow's own business
        y = (x.astype(np.float64) @ W.astype(np.float64) + b.astype(np.float64)).astype(np.float32)
        if self.activation == "relu":
            y = np.maximum(y, np.float32(0))
        elif self.activation is not None:
            raise NotImplementedError(self.activation)
        out = tf.Tensor(y)
        if ad.recording(inputs, self.kernel, self.bias):
            relu = self.activation == "relu"

            def vjp(g):
                g = (g * (y > 0)) if relu else g
                g64 = g.astype(np.float64)
                gx = (g64 @ W.astype(np.float64).T).astype(np.float32) if ad.tracked(inputs) else None
                return gx, (x.astype(np.float64).T @ g64).astype(np.float32), g64.sum(0).astype(np.float32)
            ad.attach(out, (inputs, self.kernel, self.bias), vjp)
        return out


class LeakyReLU(Layer):
    def __init__(self, alpha=0.3, **kwargs):
        super().__init__()
        self.alpha = np.float32(alpha)

    def call(self, inputs):
        import tensorflow as tf
        a = inputs.numpy()
        return tf.Tensor(np.where(a > 0, a, a * self.alpha))
