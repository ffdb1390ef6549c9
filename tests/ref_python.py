"""Helper of tests/test_ref_python.py: import the reference's OWN Python (network/*.py, util/*.py from /root/reference)
over the NumPy stand-in for TensorFlow in oracle/tf_shim (TEST INFRASTRUCTURE ONLY)."""

import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")


def available() -> bool:
    return (REF / "network" / "renderer.py").exists()


def load():
    """(tf stand-in, reference `network.renderer`, `network.model`, `network.loss`, `network.ray_sampler`, `network.proxy`,
    `util.util`).  Refuses to run next to a real TensorFlow."""
    shim = str(ROOT / "oracle" / "tf_shim")
    for p in (str(REF), shim):
        if p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, str(REF))
    sys.path.insert(0, shim)
    import tensorflow as tf
    assert "numpy-stand-in" in tf.__version__, "a real TensorFlow is importable: use it instead of the stand-in"
    from network import loss, model, proxy, ray_sampler, renderer
    from util import util
    return tf, renderer, model, loss, ray_sampler, proxy, util


def ff(util, n):
    return util.EasyDict({"module": "network.layer.FourierFeatures", "n_freq_bands": n})


def set_weights(keras_model, oracle_model):
    """Copy the oracle model's weights into the reference's Keras model: Dense layers in creation = graph order."""
    dense = [l for l in keras_model.layers if type(l).__name__ == "Dense"]
    assert len(dense) == len(oracle_model.layers)
    for l, (_, W, b) in zip(dense, oracle_model.layers):
        assert l.units == W.shape[1]
        l.set_weights([W, b])


def build_param_nerf(model_mod, util, oracle_model, name="model"):
    om = oracle_model
    m = model_mod.ParamNerf(ff(util, om.pos_freqs), ff(util, om.dir_freqs), ff(util, om.param_freqs), n_parameters=list(om.n_parameters),
                            depth=om.depth, width=om.width, skips=list(om.skips), color_depth=om.color_depth, name=name)[name]
    set_weights(m, om)
    return m
