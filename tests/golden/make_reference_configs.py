"""Extract the hot-path sub-dicts of the reference's shipped configs into a fixture ("configs as
fixtures: accept these dicts' model_config / renderer_config sub-dicts unchanged").

Run in the build container, where /root/reference exists:  python tests/golden/make_reference_configs.py
The config files are plain Python dict literals (no imports); each is executed and the keys the hot path consumes are
copied verbatim into tests/golden/reference_configs.json.  The GPU box has no /root/reference, so the GPU tests read the
JSON; tests/test_reference_configs.py re-derives it on the CPU side whenever the reference tree is present."""

import glob
import importlib.util
import json
import os
import sys

KEEP = ("module", "model_config", "renderer_config", "loss_config", "n_iters", "lrate", "lrate_decay", "seed")


def extract(reference_root: str) -> dict:
    out = {}
    for path in sorted(glob.glob(os.path.join(reference_root, "configs", "config_*.py"))):
        spec = importlib.util.spec_from_file_location("_ref_config", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        out[os.path.basename(path)[:-3]] = {k: mod.config[k] for k in KEEP if k in mod.config}
    return out


if __name__ == "__main__":
    root = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "reference_configs.json"), "w") as f:
        json.dump(extract(root), f, indent=1, sort_keys=True)
        f.write("\n")
