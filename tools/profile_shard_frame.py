"""Where one rank's time goes when an 800x800 frame is split over `world` GPUs (device-instancer pipeline of bench.py):
emulates rank 0's share on one GPU without the collective, phase by phase (CUDA events, a synchronize between phases),
then the un-synchronised wall time per frame (host-bound or not)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from nerf_tex_b200.image import finish_image  # noqa: E402
from nerf_tex_b200.sharded import shard_plan  # noqa: E402

if __name__ == "__main__":
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    dev = torch.device("cuda:0")
    pipe = bench.DeviceFramePipeline(dev, 1)
    H, W = bench.H, bench.W
    per, idx_all, restore = shard_plan(H * W, world, pipe.STRIPE, True, dev)
    idx = idx_all[0]
    params = pipe.params_host.to(dev)
    ids = idx.to(torch.int32)

    def phases(pose, record):
        def mark(name):
            if record is not None:
                torch.cuda.synchronize()
                record.append((name, time.perf_counter()))
        mark("start")
        rays_o, rays_d, t, cone = pipe.sampler(None, pose)
        mark("rays")
        mine = [a[idx].contiguous() for a in (rays_o, rays_d, t, cone)]
        mark("stripes")
        out = pipe.r(mine[0][None], mine[1][None], mine[2][None], params, mine[3][None], composite_bkgd=True, training=False, ray_ids=ids)
        mark("render")
        rec = torch.zeros(per * 4 + 1, device=dev)
        body = rec[:-1].view(per, 1, 4)
        body[:idx.shape[0], :, :3] = out["color_pred"].reshape(1, -1, 3).transpose(0, 1)
        body[:idx.shape[0], :, 3:] = out["alpha_pred"].reshape(1, -1, 1).transpose(0, 1)
        gathered = rec[:-1].repeat(world).view(world * per, 1, 4)                 # stands in for the all-gather
        full = restore(gathered)
        color, alpha = full[:, :, :3].transpose(0, 1).contiguous(), full[:, :, 3].transpose(0, 1).contiguous()
        mark("pack/unpack")
        _, u8 = finish_image(color, alpha, H, W, 1, straight_alpha=True, want_u8=True)
        pipe.host_img.copy_(u8, non_blocking=True)
        mark("finish+d2h")

    for i in range(3):
        phases(pipe.poses[i % 5], None)
    rec = []
    phases(pipe.poses[0], rec)
    out = {"world": world, "rays_mine": int(idx.shape[0])}
    for (n0, t0), (n1, t1) in zip(rec[:-1], rec[1:]):
        out[n1 + "_ms"] = (t1 - t0) * 1e3
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(10):
        phases(pipe.poses[i % 5], None)
    torch.cuda.synchronize()
    out["frame_ms_unsynced"] = (time.perf_counter() - t0) / 10 * 1e3
    print(json.dumps(out, indent=1))
