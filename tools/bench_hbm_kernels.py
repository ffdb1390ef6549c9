"""K2 / K3 micro-benchmark alone (the `hbm_kernels` block of bench.py) for quick kernel iteration."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(bench.hbm_kernels_bench(torch.device("cuda:0"), bench.peaks()), indent=1))
