"""Stand-ins for the two collaborators of the reference's network/train.py::Train that are out of scope here (SURVEY.md
section 2): the tf.data dataset and the TensorBoard / checkpoint logger.  Instantiated by the reference's own
util.instantiate from the config dicts tests/test_ref_python.py passes (TEST INFRASTRUCTURE ONLY)."""

LAST = {}


class Batches:
    """network.dataset.Dataset as far as Train uses it: n_parameters, composite_bkgd, bkgd_color, take(n) -> batches."""

    def __init__(self, batches=(), n_parameters=None, composite_bkgd=True, bkgd_color=(1., 1., 1.), step=None, **kwargs):
        self.batches, self.n_parameters = list(batches), n_parameters
        self.composite_bkgd, self.bkgd_color = composite_bkgd, list(bkgd_color)
        self.step = step

    def take(self, n):
        return iter(self.batches[:max(int(n), 0)])


class Logger:
    """network.logger.Logger as far as Train uses it: `step` (restored from a checkpoint: here 0, and the "checkpoint" is
    the set of initial weights the test wants), called once per iteration with the loss."""

    def __init__(self, target_path=None, checkpoint_variables=None, dataset=None, renderer=None, n_iters=None, initial=None, **kwargs):
        self.step, self.losses = 0, []
        self.checkpoint_variables, self.renderer = checkpoint_variables, renderer
        if initial is not None:                                       # {model name: [(kernel, bias), ...] in graph order}
            for name, layers in initial.items():
                dense = [l for l in checkpoint_variables[name].layers if type(l).__name__ == "Dense"]
                assert len(dense) == len(layers)
                for l, (W, b) in zip(dense, layers):
                    l.set_weights([W, b])
        LAST["logger"] = self

    def __call__(self, scalars):
        self.losses.append(float(scalars["Loss"]))
