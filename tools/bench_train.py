"""The `train_step` block of bench.py alone."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.train_step_bench(torch.device("cuda:0"), steps=10), indent=1))
