"""The device-instancer frame pipeline of bench.py (its `e2e` arm) alone, one GPU: pose in, 8-bit frame out."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    pipe = bench.DeviceFramePipeline(dev, 1)
    for i in range(3):
        pipe.frame(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(5):
        pipe.frame(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    out = pipe.workload_stats()
    out.update({"ms_per_frame": ms, "rays_per_sec": bench.H * bench.W / ms * 1e3, "instancer_ms_per_chunk": pipe.instancer_chunk_ms()})
    print(json.dumps(out, indent=1))
