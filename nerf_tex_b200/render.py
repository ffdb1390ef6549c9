"""Render out images (network/render.py:6-29)."""

from __future__ import annotations

from . import util


def Render(target_path: str, test_dataset_config, model_config, renderer_config, logger_config, source_path: str = None,
           override: bool = True, **kwargs):
    """Same flow as the reference: dataset -> model (n_parameters defaulting to the dataset's) -> renderer ->
    logger, which restores the latest checkpoint and renders every dataset entry to `target_path`/media/test.
    ``test_dataset_config`` may be a config dict (e.g. module ``network.dataset.PoseDataset``) or a ready dataset
    object; the tf.data based loaders of network/dataset.py are out of scope.  Returns the logger."""
    test_dataset = util.instantiate(test_dataset_config) if isinstance(test_dataset_config, dict) else test_dataset_config
    model_config = util.EasyDict(model_config)
    model_config.setdefault("n_parameters", test_dataset.n_parameters)                  # :20
    model = util.instantiate(model_config)
    renderer_config = util.EasyDict(renderer_config)
    renderer_config.update(model)                                                       # :24
    renderer = util.instantiate(renderer_config)
    logger_config = util.EasyDict(logger_config)
    logger_config.update({"target_path": target_path, "checkpoint_variables": dict(model), "source_path": source_path,
                          "dataset": test_dataset, "is_training": False, "renderer": renderer})   # :28
    return util.instantiate(logger_config)
