"""The `device_instancer_frame` block of bench.py alone."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.device_instancer_bench(torch.device("cuda:0")), indent=1))
