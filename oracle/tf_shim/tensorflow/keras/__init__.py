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
