"""tensorflow.keras of the NumPy stand-in (TEST INFRASTRUCTURE ONLY): the functional API as far as the reference's
network/model.py uses it — Input placeholders, layers called on them, slicing / tf.concat of symbolic tensors, Model —
evaluated lazily when the Model is called on real tensors."""

from __future__ import annotations

import numpy as np

from . import layers  # noqa: F401
from .layers import KerasTensor, resolve


def Input(shape=None, name=None, dtype=np.float32, **kwargs):
    return KerasTensor(None, name=name, shape=shape)


class Model:
    def __init__(self, inputs=None, outputs=None, name=None):
        self.inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        self.outputs = outputs
        self.name = name
        # every layer created while this model's graph was written, in creation order (Dense layers: graph order)
        self.layers = layers.collect_layers(outputs)

    def __call__(self, inputs, training=False):
        inputs = list(inputs) if isinstance(inputs, (list, tuple)) else [inputs]
        env = {id(p): v for p, v in zip(self.inputs, inputs)}
        cache = {}
        if isinstance(self.outputs, (list, tuple)):
            return [resolve(o, env, cache) for o in self.outputs]
        return resolve(self.outputs, env, cache)

    @property
    def trainable_variables(self):
        out = []
        for l in self.layers:
            out += l.trainable_variables
        return out

    def summary(self, *a, **k):
        pass


class utils:
    @staticmethod
    def plot_model(*args, **kwargs):
        pass


class optimizers:
    """tf.keras.optimizers as documented for TF 2.4: Adam (non-amsgrad, epsilon inside the denominator AFTER the square
    root, bias correction folded into the step size) and the ExponentialDecay schedule."""

    class schedules:
        class ExponentialDecay:
            def __init__(self, initial_learning_rate, decay_steps, decay_rate, staircase=False, name=None):
                self.initial_learning_rate, self.decay_steps, self.decay_rate, self.staircase = initial_learning_rate, decay_steps, decay_rate, staircase

            def __call__(self, step):
                p = float(step) / float(self.decay_steps)
                if self.staircase:
                    p = np.floor(p)
                return np.float32(self.initial_learning_rate) * np.float32(np.power(np.float32(self.decay_rate), np.float32(p)))

    class Adam:
        def __init__(self, learning_rate=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, amsgrad=False, name="Adam"):
            assert not amsgrad
            self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
            self.iterations = 0
            self._m, self._v = {}, {}

        def _lr(self):
            lr = self.learning_rate
            return float(lr(self.iterations)) if callable(lr) else float(lr)

        def apply_gradients(self, grads_and_vars):
            lr = self._lr()
            t = self.iterations + 1
            lr_t = lr * np.sqrt(1.0 - self.beta_2 ** t) / (1.0 - self.beta_1 ** t)
            for g, v in grads_and_vars:
                if g is None:
                    continue
                g = np.asarray(g.numpy() if hasattr(g, "numpy") else g, np.float32)
                m = self._m.get(id(v), np.zeros_like(g))
                s = self._v.get(id(v), np.zeros_like(g))
                m = (self.beta_1 * m + (1.0 - self.beta_1) * g).astype(np.float32)
                s = (self.beta_2 * s + (1.0 - self.beta_2) * g * g).astype(np.float32)
                self._m[id(v)], self._v[id(v)] = m, s
                v.assign_sub((lr_t * m / (np.sqrt(s) + self.epsilon)).astype(np.float32))
            self.iterations = t

