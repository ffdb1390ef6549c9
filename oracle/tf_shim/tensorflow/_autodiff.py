"""Reverse-mode differentiation for the NumPy stand-in (TEST INFRASTRUCTURE ONLY): enough of tf.GradientTape to run the
training step of the reference's network/train.py:60-67 — the operations on the path from the Dense variables through
Renderer.map_model_output, Renderer.__call__'s scatter and the losses of network/loss.py.  An operation without a
rule that receives a tracked tensor yields an untracked result and is noted in `DROPPED` (the test checks that only the
operations behind tf.stop_gradient land there)."""

from __future__ import annotations

import numpy as np

TAPES = []
DROPPED = set()


class Node:
    __slots__ = ("parents", "vjp")

    def __init__(self, parents, vjp):
        self.parents, self.vjp = parents, vjp


LEAF = Node((), None)


def tracked(x):
    return getattr(x, "_node", None) is not None


def recording(*xs):
    return bool(TAPES) and any(tracked(x) for x in xs)


def attach(out, parents, vjp):
    """`vjp(g) -> one gradient (ndarray or None) per parent`."""
    out._node = Node(tuple(parents), vjp)
    return out


def dropped(name, *xs):
    if recording(*xs):
        DROPPED.add(name)


def unbroadcast(g, shape):
    g = np.asarray(g)
    while g.ndim > len(shape):
        g = g.sum(0)
    for ax, n in enumerate(shape):
        if n == 1 and g.shape[ax] != 1:
            g = g.sum(ax, keepdims=True)
    return g.reshape(shape)


def backward(loss, sources):
    order, seen = [], set()

    def visit(t):
        if id(t) in seen or not tracked(t):
            return
        seen.add(id(t))
        for p in t._node.parents:
            visit(p)
        order.append(t)
    import sys
    sys.setrecursionlimit(max(sys.getrecursionlimit(), 100000))
    visit(loss)
    grads = {id(loss): np.ones_like(loss._a)}
    for t in reversed(order):
        g = grads.get(id(t))
        if g is None or t._node.vjp is None:
            continue
        for p, gp in zip(t._node.parents, t._node.vjp(g)):
            if gp is None or not tracked(p):
                continue
            gp = np.asarray(gp, p._a.dtype if p._a.dtype.kind == "f" else np.float32)
            grads[id(p)] = gp if id(p) not in grads else grads[id(p)] + gp
    return [grads.get(id(s)) for s in sources]


class GradientTape:
    def __init__(self, persistent=False, watch_accessed_variables=True):
        pass

    def __enter__(self):
        TAPES.append(self)
        return self

    def __exit__(self, *exc):
        TAPES.remove(self)
        return False

    def watch(self, t):
        if not tracked(t):
            t._node = LEAF

    def gradient(self, target, sources):
        import tensorflow as tf
        single = not isinstance(sources, (list, tuple))
        gs = backward(target, [sources] if single else list(sources))
        gs = [None if g is None else tf.Tensor(g) for g in gs]
        return gs[0] if single else gs
