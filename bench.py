#!/usr/bin/env python
"""bench.py — headline benchmark of the NeRF-Tex hot path on B200.

Workload (BASELINE.json configs[1], SURVEY.md section 8d "C2"): one 800x800 frame (640 000 rays) of the
carpet-shape instanced render — model n_parameters [1,6], InstanceRenderer kwargs of
configs/config_carpet_render.py:76-99 (n_samples 1024, step 0.002, patch_scale 0.09, render_chunk
16384), random-init weights, seeded synthetic instancer output (hit ~ 0.7, in-shell samples per
hit ray ~ LogNormal(ln 120, 0.6)).  A "step" is one frame: per render chunk one fused
encoding+MLP launch (K1) and one compositing launch (K2).

  value : rays/s, whole job, instancer records resident in HBM when the timed region starts.  For N > 1 the SAME
          frame is split over the ranks (round-robin stripes of 256 rays, sharded.render_sharded) with the final
          all-gather of the per-ray outputs inside the timed region ("scaling": "strong").
  e2e   : rays/s through the call a user makes for a render with the device instancer: camera pose + parameter vector
          in from the host (92 B), Proxy ray sampler -> InstanceRenderer(Instancer(1 500-patch carpet)) -> finish_image,
          finished 8-bit RGBA frame out to pinned host memory (2.56 MB), every frame; sharded the same way for N > 1.
          (Round 1's arm — host-side instancer records copied per chunk — is kept at N = 1 as `e2e_host_records`.)
  --impl reference : the reference's CPU implementation of the same path.  The reference's
          TensorFlow cannot be installed here (and /root/reference is not on the GPU box), so this arm times the
          oracle port (oracle/torch_port.py, torch-CPU fp32, all host threads) on a bounded ray subset.

One process per GPU; for N > 1 launch with torch.distributed.run.  Rank 0 prints ONE JSON line; N > 1 lines carry
`checks` (sharded frame == single-GPU frame, all-reduced gradient == mean, replicas identical).  Both timed regions are
bracketed by NVTX ranges (ntx_value_timed, ntx_e2e_timed) for `ncu --nvtx --nvtx-include`.
"""

from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

H = W = 800
RENDER_CHUNK = 16384
N_SAMPLES = 1024
STEP_SIZE = 0.002
PATCH_SCALE = 0.09
NPAR = (1, 6)
B0, B1 = [-1.4, -1.2, -.1], [1.2, 1.2, 1.8]
DATASET_PARAMS = [1, 1, 1, .1, 0, 0, 1]          # config_carpet_render.py:40
CPU_SAMPLE_RAYS = 8192
MACS_PER_SAMPLE = {(1, 6): 680832, (1, 4): 676224, (2, 3): 678528}


# dram__bytes_read.sum + dram__bytes_write.sum of one mlp_forward_kernel chunk launch (profiles/r01_v6_k1_full_raw.csv)
K1_NCU_DRAM_BYTES_PER_LAUNCH = 86104320 + 10907136


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def frame_rays(seed: int):
    """800x800 camera rays (ray_sampler.rays_from_camera semantics) with a constant near/far."""
    rng = np.random.default_rng(seed)
    focal = 0.5 * W / math.tan(0.5 * 0.63)
    ii, jj = np.meshgrid(np.arange(H, dtype=np.float32), np.arange(W, dtype=np.float32), indexing="ij")
    dirs = np.stack([(jj + .5 - .5 * W) / focal, -(ii + .5 - .5 * H) / focal, -np.ones_like(ii)], -1).reshape(-1, 3)
    ang = rng.uniform(0, 2 * math.pi)
    c2w = np.array([[math.cos(ang), 0, math.sin(ang)], [0, 1, 0], [-math.sin(ang), 0, math.cos(ang)]], np.float32)
    rays_d = (dirs @ c2w.T).astype(np.float32)
    rays_d /= np.linalg.norm(rays_d, axis=-1, keepdims=True)
    rays_o = np.tile((c2w @ np.array([0, 0, 4.], np.float32))[None], (H * W, 1)).astype(np.float32)
    t = np.tile(np.array([[2., 6.]], np.float32), (H * W, 1))
    cone = (np.cos(np.arctan(np.linalg.norm(dirs[:, :2], axis=-1))) / np.linalg.norm(dirs, axis=-1) / focal)
    return rays_o[None], rays_d[None], t[None], np.array([DATASET_PARAMS], np.float32), cone.astype(np.float32)[None, :, None]


class ChunkReplayInstancer:
    """Instancer boundary object for the bench: replays pre-generated per-chunk CSR records.
    resident=True: records live in HBM.  resident=False: records live in pinned host memory and
    every call copies its chunk to the device (the e2e arm)."""

    patch_scale = PATCH_SCALE

    def __init__(self, chunks, resident: bool, fields=None):
        import torch
        if fields is not None:           # host side: ship only the records this renderer configuration consumes
            chunks = [{k: v for k, v in c.items() if k in fields} for c in chunks]
        self.resident = resident
        self.i = 0
        self.h2d_bytes = 0
        self.copy_stream = None
        self.pending = None
        if resident:
            self.chunks = chunks
        else:
            self.chunks = [{k: v.cpu().pin_memory() for k, v in c.items()} for c in chunks]

    def n_instances(self):
        return 10000

    def reset(self):
        self.i = 0

    def _start_copy(self, idx, device):
        """Enqueue the host->device copy of chunk `idx` on the side stream."""
        import torch
        if self.copy_stream is None:
            self.copy_stream = torch.cuda.Stream(device=device)
        with torch.cuda.stream(self.copy_stream):
            out = {}
            for k, v in self.chunks[idx].items():
                out[k] = v.to(device, non_blocking=True)
                self.h2d_bytes += v.numel() * v.element_size()
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return out, ev

    def get_model_input_csr(self, rays_o, rays_d, parameters, n_samples, step_size):
        import torch
        idx = self.i
        self.i = (self.i + 1) % len(self.chunks)
        if self.resident:
            return self.chunks[idx]
        # host instancer: records arrive from pinned host memory; the copy of the NEXT chunk is
        # started before this one is consumed so that PCIe transfers overlap the kernels
        if self.pending is None or self.pending[0] != idx:
            self.pending = (idx,) + self._start_copy(idx, rays_o.device)
        _, cur, ev = self.pending
        nxt = (idx + 1) % len(self.chunks)
        self.pending = (nxt,) + self._start_copy(nxt, rays_o.device)
        torch.cuda.current_stream().wait_event(ev)
        for v in cur.values():
            v.record_stream(torch.cuda.current_stream())
        return cur


def clocks_start():
    path = Path(tempfile.gettempdir()) / f"ntx_clocks_{os.getpid()}.csv"
    try:
        proc = subprocess.Popen(
            ["nvidia-smi", "-i", os.environ.get("LOCAL_RANK", "0"),
             "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"],
            stdout=open(path, "w"), stderr=subprocess.DEVNULL)
        return proc, path
    except Exception:
        return None, path


def clocks_stop(proc, path):
    if proc is None:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for line in Path(path).read_text().splitlines():
        f = [x.strip() for x in line.split(",")]
        if len(f) < 7:
            continue
        try:
            sm.append(float(f[0])); mx.append(float(f[1]))
        except ValueError:
            continue
        for n, v in zip(names, f[3:7]):
            if v.lower().startswith("active"):
                reasons.add(n)
    # samples under load = the upper half of the observed clocks (idle samples bracket the region)
    load = sorted(sm)[len(sm) // 2:] if sm else []
    return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
            "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline_run(steps: int, warmup: int):
    """Oracle port of InstanceRenderer.evaluate_model (renderer.py:247-354) on CPU_SAMPLE_RAYS rays."""
    from oracle import nerftex_oracle as O
    rng = np.random.default_rng(0)
    R, S, p = CPU_SAMPLE_RAYS, N_SAMPLES, sum(NPAR)
    hit = rng.uniform(size=R) < 0.7
    idxs = np.nonzero(hit)[0]
    n_valid = np.clip(np.round(np.exp(rng.normal(size=R) * 0.6 + math.log(120.))), 1, S).astype(np.int64)
    n_valid[~hit] = 0
    mask = np.arange(S)[None, :] < n_valid[:, None]
    dists = np.where(mask, np.float32(STEP_SIZE), np.float32(0)).astype(np.float32)
    pts = rng.uniform(B0, B1, (R, S, 3)).astype(np.float32)
    d = rng.normal(size=(R, S, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    params = np.tile(np.array(DATASET_PARAMS, np.float32), (R, S, 1)) * rng.uniform(0.5, 1, (R, S, p)).astype(np.float32)
    dense = (d, pts, np.zeros((R, S), np.float32), dists, np.zeros((R, 1, 3), np.float32),
             (rng.uniform(size=(R, 1)) < .5).astype(np.float32), np.ones((R, S), np.float32),
             np.zeros((R, S), np.int32), idxs, params)
    from oracle.torch_port import TorchPort
    model = O.ParamNerfOracle(n_parameters=NPAR, seed=1)
    port = TorchPort(model)                                   # threaded fp32 port of the same graph (pinned to the
    n_in_shell = int(mask.sum())                              # NumPy oracle in tests/test_torch_port.py)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        port.instance_evaluate_model(dense, R, patch_scale=PATCH_SCALE, net_chunk=32768)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    return {"value": R / dt, "unit": "rays/s", "cores": cores, "kind": "port",
            "sample": f"{R} rays of the frame ({n_in_shell} in-shell samples), oracle/torch_port.py "
                      f"instance_evaluate_model (torch-CPU fp32, all host threads), mean of {steps} passes",
            "samples_per_sec": n_in_shell / dt, "ms_per_pass": dt * 1e3}


def hbm_kernels_bench(dev, pk):
    """K2 / K3 alone on inputs far larger than L2: achieved GB/s from ALGORITHMIC bytes
     over CUDA-event time, against the measured HBM copy bandwidth."""
    import torch
    from nerf_tex_b200 import ops
    out = {}
    g = torch.Generator(device=dev).manual_seed(0)

    def timeit(fn, reps=5):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps / 1e3

    R, S = 1 << 20, 192                                        # 3.2 GB of rgb-sigma: >> 126 MB L2
    rgbs = torch.randn((R, S, 4), device=dev, generator=g)
    z = torch.sort(torch.rand((R, S), device=dev, generator=g) * 4 + 2, -1).values
    rays_d = torch.randn((R, 3), device=dev, generator=g)
    dt = timeit(lambda: ops.composite_forward(rgbs, z, rays_d, want_weights=True))
    algo = R * (24 * S + 28)
    out["composite_forward"] = {"rays": R, "samples_per_ray": S, "bytes_per_ray": 24 * S + 28, "ms": dt * 1e3,
                                "achieved_gbs": algo / dt / 1e9, "frac_of_measured_hbm": algo / dt / 1e9 / pk["hbm_gbs"],
                                "rays_per_sec": R / dt}
    del rgbs
    # the instanced compositor (the K2 variant the headline workload launches): ragged CSR rows, ~144 samples per ray
    Hn = 1 << 17
    cnt = torch.clamp(torch.round(torch.exp(torch.randn(Hn, device=dev, generator=g) * 0.6 + math.log(120.0))), 1, 1024).long()
    ptr = torch.zeros(Hn + 1, dtype=torch.int64, device=dev)
    ptr[1:] = torch.cumsum(cnt, 0)
    T = int(ptr[-1])
    rg = torch.randn((T, 4), device=dev, generator=g)
    di = torch.full((T,), 0.002, device=dev)
    aw = torch.ones(T, device=dev)
    cl, al = torch.zeros((Hn, 3), device=dev), (torch.rand(Hn, device=dev, generator=g) < .5).float()
    ri = torch.arange(Hn, device=dev, dtype=torch.int32)
    dt = timeit(lambda: ops.composite_instanced_forward(rg, di, aw, ptr, cl, al, ri, Hn, patch_scale=0.09))
    algo = 24 * T + 32 * Hn
    out["composite_instanced_forward"] = {"rays": Hn, "samples": T, "bytes": algo, "ms": dt * 1e3, "achieved_gbs": algo / dt / 1e9,
                                          "frac_of_measured_hbm": algo / dt / 1e9 / pk["hbm_gbs"], "rays_per_sec": Hn / dt}
    del rg, di, aw
    S0, N = 64, 128
    R3 = 1 << 21
    z0 = torch.sort(torch.rand((R3, S0), device=dev, generator=g) * 4 + 2, -1).values
    w0 = torch.rand((R3, S0), device=dev, generator=g) ** 8
    dt = timeit(lambda: ops.hierarchical_z(z0, w0, N, True))
    algo = R3 * (S0 * 4 + S0 * 4 + (S0 + N) * 4)
    out["hierarchical_z"] = {"rays": R3, "bytes_per_ray": S0 * 8 + (S0 + N) * 4, "ms": dt * 1e3,
                             "achieved_gbs": algo / dt / 1e9, "frac_of_measured_hbm": algo / dt / 1e9 / pk["hbm_gbs"],
                             "rays_per_sec": R3 / dt}
    return out


def train_step_bench(dev, steps=5):
    """SURVEY.md C3 (grass_filtered shape): 4096 rays, model [2,3], blur_idx 0, raw_noise_std 0.1,
    AlphaLoss(smape, mse); variant (i) single net x 256 samples and (ii) CoarseFine 64 + 128.
    Whole step: forward (K3/K1/K2), loss, backward (K2-bwd + MLP dgrad/wgrad), Adam."""
    import torch
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.train import Trainer
    g = torch.Generator(device=dev).manual_seed(0)
    R = 4096
    ro = torch.tensor([[0., 0.5, 5.]], device=dev).repeat(R, 1)
    rd = torch.rand((R, 3), device=dev, generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    data = {"rays_o": ro[None], "rays_d": rd[None],
            "t": torch.stack([3 + torch.rand(R, device=dev, generator=g), 6 + torch.rand(R, device=dev, generator=g)], -1)[None],
            "parameters": torch.rand((1, 5), device=dev, generator=g), "cone_scale": torch.rand((1, R, 1), device=dev, generator=g) * .01,
            "color": torch.rand((1, R, 3), device=dev, generator=g), "alpha": torch.rand((1, R), device=dev, generator=g)}
    out = {}
    for name, (S, N), fp8 in (("single_256", (256, 0), False), ("coarse_fine_64_128", (64, 128), False),
                              ("single_256_fp8_saved", (256, 0), True), ("coarse_fine_64_128_fp8_saved", (64, 128), True)):
        models = {"model": FusedNerfTexMLP(2, 3, seed=1)}
        if N:
            models["model_fine"] = FusedNerfTexMLP(2, 3, seed=2, name="model_fine")
        if fp8:                                    # saved tile images in fp8 e4m3 (first step calibrates in bf16)
            for m in models.values():
                m.enable_fp8_saved_operands()
        r = Renderer(n_samples=S, n_importance=N, perturb=True, blur_idx=0, raw_noise_std=0.1, seed=0, **models)
        tr = Trainer(models, r, AlphaLoss("network.loss.smape", "network.loss.mse"))
        for _ in range(3):
            tr.train_step(data)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = tr.train_step(data)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        rows = R * 256
        out[name] = {"ms_per_step": ms, "rays_per_sec": R / ms * 1e3, "mlp_rows_per_step": rows,
                     "fwd_bwd_tflops": rows * 2.92 * 2 * 678528 / (ms / 1e3) / 1e12, "loss": float(loss)}
    return out


def train_step_dp_bench(dev, rank: int, world: int, steps: int = 10):
    """BASELINE.json configs[3] (fur training shape): data-parallel training, model [1,4], 4096 rays x
    256 samples per rank (seed 1000 + rank), identical initial weights, one NCCL all-reduce of the flat fp32 gradient
    buffer per step, Adam on every rank.  Step time = max over ranks; throughput = rays of all ranks / that time."""
    import torch
    import torch.distributed as dist
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.train import Trainer
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    R = 4096
    ro = torch.tensor([[0., 0.5, 5.]], device=dev).repeat(R, 1)
    rd = torch.rand((R, 3), device=dev, generator=g) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    data = {"rays_o": ro[None], "rays_d": rd[None],
            "t": torch.stack([3 + torch.rand(R, device=dev, generator=g), 6 + torch.rand(R, device=dev, generator=g)], -1)[None],
            "parameters": torch.rand((1, 5), device=dev, generator=g), "cone_scale": torch.rand((1, R, 1), device=dev, generator=g) * .01,
            "color": torch.rand((1, R, 3), device=dev, generator=g), "alpha": torch.rand((1, R), device=dev, generator=g)}
    models = {"model": FusedNerfTexMLP(1, 4, seed=1)}
    r = Renderer(n_samples=256, n_importance=0, perturb=True, seed=rank, **models)
    tr = Trainer(models, r, AlphaLoss("network.loss.smape", "network.loss.mse"), lrate=5e-4)
    # check: the all-reduced gradient is the mean of the per-rank gradients (gathered separately)
    allreduce_eq_mean = True
    if world > 1:
        pred = r(**{k: data[k] for k in ("rays_o", "rays_d", "t", "parameters", "cone_scale")}, training=True)
        tr.loss_fn(color_true=data["color"], alpha_true=data["alpha"], **pred).backward()
        m = models["model"]
        local_g = m.master.grad.detach().clone()
        gathered = [torch.empty_like(local_g) for _ in range(world)]
        dist.all_gather(gathered, local_g)
        mean = torch.stack(gathered).double().sum(0) / world
        differ = bool((gathered[0] != gathered[-1]).any())
        tr.allreduce_gradients()
        err = float((m.master.grad.double() - mean).abs().max() / (mean.abs().max() + 1e-30))
        flags = [None] * world
        dist.all_gather_object(flags, bool(err < 1e-6 and differ))
        allreduce_eq_mean = all(flags)
        m.master.grad = None
    for _ in range(3):
        tr.train_step(data)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = tr.train_step(data)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    e0.record()
    for _ in range(20):
        tr.allreduce_gradients()
    e1.record()
    torch.cuda.synchronize()
    ar = torch.tensor([e0.elapsed_time(e1) / 20 * 1e3], device=dev)
    chk = models["model"].master.detach().double().sum().reshape(1)          # replicas must stay identical
    lo, hi = chk.clone(), chk.clone()
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(ar, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    return {"model": "[1,4]", "rays_per_rank": R, "samples_per_ray": 256, "ms_per_step": float(ms), "n_gpus": world,
            "rays_per_sec": world * R / float(ms) * 1e3, "allreduce_us": float(ar), "grad_bytes": int(models["model"].n_params * 4),
            "loss_rank0": float(loss),
            "checks": {"allreduce_eq_mean": bool(allreduce_eq_mean), "replicas_identical": bool(float(lo) == float(hi))}}


def c1_forward_bench(dev, steps=20):
    """BASELINE.json configs[0]: 4096 rays x (64 coarse + 128 fine), CoarseFine [1,4], forward only,
    through Renderer.__call__ (K3 -> K1 -> K2 -> K3 -> K1 -> K2)."""
    import torch
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    g = torch.Generator(device=dev).manual_seed(0)
    B, Nr = 16, 256
    o = torch.randn((B, 3), device=dev, generator=g)
    o[:, 2] = o[:, 2].abs()
    o = 5 * o / o.norm(dim=-1, keepdim=True)
    rays_o = o[:, None, :].repeat(1, Nr, 1)
    rays_d = torch.rand((B, Nr, 3), device=dev, generator=g) * 2 - 1 - rays_o
    rays_d = rays_d / rays_d.norm(dim=-1, keepdim=True)
    t = torch.stack([3.5 + torch.rand((B, Nr), device=dev, generator=g) * .5, 5.5 + torch.rand((B, Nr), device=dev, generator=g)], -1)
    params = torch.rand((B, 5), device=dev, generator=g)
    cone = torch.full((B, Nr, 1), 2e-3, device=dev)
    r = Renderer(model=FusedNerfTexMLP(1, 4, seed=1), model_fine=FusedNerfTexMLP(1, 4, seed=2, name="model_fine"),
                 n_samples=64, n_importance=128, perturb=True, seed=0)
    for _ in range(5):
        r(rays_o, rays_d, t, params, cone, training=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r(rays_o, rays_d, t, params, cone, training=False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    rows = B * Nr * 256
    return {"rays": B * Nr, "mlp_rows": rows, "ms_per_batch": ms, "rays_per_sec": B * Nr / ms * 1e3,
            "samples_per_sec": rows / ms * 1e3, "mlp_tflops_incl_everything": rows * 2 * 676224 / (ms / 1e3) / 1e12,
            "note": "4096-ray batch fills 148 SMs for only ~7 waves of 512-row CTA-pair batches; launch-latency sensitive"}


def c5_plush_bench(dev, steps=3):
    """BASELINE.json configs[4] / SURVEY.md C5 on one GPU: 800x800 frame, model [1,4], InstanceRenderer kwargs of
    configs/config_plush_render.py:87-112 (n_samples 1280, step 0.0005, patch_scale 0.04, render_chunk 32768), seeded
    synthetic instancer with ~LogNormal(ln 200, 0.6) in-shell samples per hit ray, records resident in HBM."""
    import torch
    from nerf_tex_b200.instancer import SyntheticInstancer
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import InstanceRenderer
    chunk, S, step, ps = 32768, 1280, 0.0005, 0.04
    rays_o, rays_d, t, _, cone = frame_rays(seed=1005)
    params = np.array([[1, 1, 0, 0, 1]], np.float32)
    synth = SyntheticInstancer(b_0=[-1.1, -1.1, -.2], b_1=[1.1, 1.1, 1.1], patch_scale=ps, seed=5, mean_valid=200.0)
    d_o, d_d = torch.as_tensor(rays_o[0]).to(dev), torch.as_tensor(rays_d[0]).to(dev)
    d_p = torch.as_tensor(params).to(dev).repeat_interleave(H * W, dim=0)
    chunks = [synth.get_model_input_csr(d_o[i:i + chunk], d_d[i:i + chunk], d_p[i:i + chunk], S, step) for i in range(0, H * W, chunk)]
    n_samples = int(sum(c["pts"].shape[0] for c in chunks))
    inst = ChunkReplayInstancer(chunks, resident=True)
    inst.patch_scale = ps
    r = InstanceRenderer(instancer_config=inst, model=FusedNerfTexMLP(1, 4, seed=1, device=dev), n_samples=S, render_chunk=chunk,
                         net_chunk=65536, step_size=step, density_reweighting=True)
    dev_in = [torch.as_tensor(x).to(dev) for x in (rays_o, rays_d, t, params, cone)]

    def frame():
        inst.reset()
        return r(*dev_in, composite_bkgd=False, training=False)
    frame()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        frame()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"rays": H * W, "in_shell_samples": n_samples, "ms_per_frame": ms, "rays_per_sec": H * W / ms * 1e3,
            "samples_per_sec": n_samples / ms * 1e3, "mlp_tflops": n_samples * 2 * MACS_PER_SAMPLE[(1, 4)] / (ms / 1e3) / 1e12}


def carpet_patches(n: int, seed: int = 0, extent: float = 1.45, patch_scale: float = 0.09):
    """Procedural stand-in for the reference's cloth mesh + anchor points (git-LFS stubs): `n` shell patches anchored
    on the z = 0 plane inside [-extent, extent]^2, each rotated about z, tilted a little and scaled by patch_scale
    (what AddInstance receives, instancer.cpp:125-141)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    T = np.tile(np.eye(4), (n, 1, 1))
    a, bx, by = rng.uniform(0, 2 * math.pi, n), rng.normal(0, 0.15, n), rng.normal(0, 0.15, n)
    ca, sa = np.cos(a), np.sin(a)
    rz = np.stack([np.stack([ca, -sa, 0 * a], -1), np.stack([sa, ca, 0 * a], -1), np.stack([0 * a, 0 * a, 1 + 0 * a], -1)], 1)
    rx = np.stack([np.stack([1 + 0 * a, 0 * a, 0 * a], -1), np.stack([0 * a, np.cos(bx), -np.sin(bx)], -1),
                   np.stack([0 * a, np.sin(bx), np.cos(bx)], -1)], 1)
    ry = np.stack([np.stack([np.cos(by), 0 * a, np.sin(by)], -1), np.stack([0 * a, 1 + 0 * a, 0 * a], -1),
                   np.stack([-np.sin(by), 0 * a, np.cos(by)], -1)], 1)
    T[:, :3, :3] = (rz @ rx @ ry) * (patch_scale * rng.uniform(0.9, 1.1, n))[:, None, None]
    T[:, 0, 3], T[:, 1, 3] = rng.uniform(-extent, extent, n), rng.uniform(-extent, extent, n)
    return T


def config_dict(extra=None):
    c = {"workload": "carpet-shape instanced render, 800x800 frame (640000 rays), n_parameters [1,6], "
                     "n_samples<=1024, step 0.002, patch_scale 0.09, render_chunk 16384, random-init weights",
         "rays_per_step": H * W, "cache": "per-step inputs (~4 GB of instancer records) exceed the 126 MB L2",
         "instancer": "synthetic (seeded): hit~0.7, in-shell samples/ray~LogNormal(ln120,0.6)"}
    if extra:
        c.update(extra)
    return c


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_baseline_run(args.steps, max(args.warmup, 1))
    line = {"impl": "reference", "metric": "rays/s (800x800 carpet-shape instanced render; samples/s alongside)",
            "value": cb["value"], "unit": "rays/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": cb["ms_per_pass"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict({"reference_note":
            "reference TensorFlow 2.4.1 is not installable in this image; oracle port timed on host cores"}),
            "samples_per_sec": cb["samples_per_sec"], "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class FrameRecords:
    """Instancer records of the whole 800x800 frame as ONE CSR over all rays (generated once by the seeded
    SyntheticInstancer in 16 384-ray calls, identical on every rank), from which any ray subset — a rank's stripes, cut
    into render chunks — is selected.  A ray's records do not depend on which rank or chunk it lands in, so the sharded
    frame can be compared bit for bit with the single-GPU one."""

    PER_SAMPLE = ("pts", "dirs", "params", "dists", "t", "alpha_weight", "instance_id")

    def __init__(self, dev, d_o, d_d, d_p):
        import torch
        from nerf_tex_b200.instancer import SyntheticInstancer
        synth = SyntheticInstancer(b_0=B0, b_1=B1, patch_scale=PATCH_SCALE, seed=0)
        n_rays = d_o.shape[0]
        parts = []
        counts = torch.zeros(n_rays, dtype=torch.int64, device=dev)
        self.hit = torch.zeros(n_rays, dtype=torch.bool, device=dev)
        self.color_last = torch.zeros((n_rays, 3), dtype=torch.float32, device=dev)
        self.alpha_last = torch.zeros(n_rays, dtype=torch.float32, device=dev)
        for i in range(0, n_rays, RENDER_CHUNK):
            c = synth.get_model_input_csr(d_o[i:i + RENDER_CHUNK], d_d[i:i + RENDER_CHUNK], d_p[i:i + RENDER_CHUNK],
                                          N_SAMPLES, STEP_SIZE)
            g = c["ray_index"].long() + i
            counts[g] = c["ray_ptr"][1:] - c["ray_ptr"][:-1]
            self.hit[g] = True
            self.color_last[g], self.alpha_last[g] = c["color_last"], c["alpha_last"]
            parts.append({k: c[k] for k in self.PER_SAMPLE})
        self.rec = {k: torch.cat([q[k] for q in parts]) for k in self.PER_SAMPLE}
        self.ptr = torch.zeros(n_rays + 1, dtype=torch.int64, device=dev)
        self.ptr[1:] = torch.cumsum(counts, 0)
        self.counts = counts
        self.n_samples_total = int(self.ptr[-1])
        self.n_hit = int(self.hit.sum())

    def chunks_for(self, ray_ids):
        """Render-chunk CSR dicts (RENDER_CHUNK rays each, in order) for the rays `ray_ids` of the frame."""
        import torch
        out = []
        for i in range(0, ray_ids.shape[0], RENDER_CHUNK):
            ids = ray_ids[i:i + RENDER_CHUNK]
            local = torch.nonzero(self.hit[ids]).squeeze(1)
            g = ids[local]
            cnt, start = self.counts[g], self.ptr[:-1][g]
            ptr = torch.zeros(g.shape[0] + 1, dtype=torch.int64, device=ids.device)
            ptr[1:] = torch.cumsum(cnt, 0)
            rows = torch.repeat_interleave(start - ptr[:-1], cnt) + torch.arange(int(ptr[-1]), device=ids.device)
            c = {k: v[rows].contiguous() for k, v in self.rec.items()}
            c.update({"ray_ptr": ptr, "ray_index": local.to(torch.int32), "color_last": self.color_last[g].contiguous(),
                      "alpha_last": self.alpha_last[g].contiguous()})
            out.append(c)
        return out


class DeviceFramePipeline:
    """The call a user makes for a render with the device instancer: camera pose + parameter vector in (host), finished
    8-bit RGBA frame out (pinned host memory).  Proxy ray sampler (ntx_rays_from_camera) -> InstanceRenderer over the
    device shell Instancer (procedural 1 500-patch carpet) -> K1 -> instanced K2 -> [all-gather of the per-ray outputs
    when the frame is sharded over ranks: round-robin stripes of 256 rays] -> ntx_image_finish -> D2H."""

    STRIPE = 256

    def __init__(self, dev, world):
        import torch
        from nerf_tex_b200.dataset import look_at
        from nerf_tex_b200.instancer import Instancer
        from nerf_tex_b200.model import FusedNerfTexMLP
        from nerf_tex_b200.proxy import AABB
        from nerf_tex_b200.ray_sampler import Proxy
        from nerf_tex_b200.renderer import InstanceRenderer
        self.dev, self.world = dev, world
        self.n_inst = 1500
        self.inst = Instancer(b_0=B0, b_1=B1, textures=["", "", "", "", "light"], transformations=carpet_patches(self.n_inst),
                              patch_scale=PATCH_SCALE, instance_sampling_method="nearest", device=dev)
        self.chunk = int(os.environ.get("NTX_BENCH_INST_CHUNK", "131072"))       # rays per instancer / K1 / K2 launch
        self.r = InstanceRenderer(instancer_config=self.inst, model=FusedNerfTexMLP(*NPAR, seed=1, device=dev),
                                  n_samples=N_SAMPLES, render_chunk=self.chunk, step_size=STEP_SIZE, density_reweighting=True)
        focal = W / math.tan(0.55 / 2) / 2
        self.sampler = Proxy(H, W, focal, AABB([-1.5, -1.5, -1.5], [1.5, 1.5, 1.5]))
        z = 1 - 2 * 0.3                                                           # data.distribution.Sphere, u = 0.3
        self.poses = [look_at([6 * math.cos(v) * math.sqrt(1 - z * z), 6 * math.sin(v) * math.sqrt(1 - z * z), 6 * z])
                      for v in (0.0, 1.3, 2.6, 3.9, 5.2)]
        self.params_host = torch.tensor([DATASET_PARAMS], dtype=torch.float32).pin_memory()
        self.host_img = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
        self.h2d_bytes = 64 + 28                                                  # 4x4 pose (by value) + 7 parameters
        self.d2h_bytes = H * W * 4

    def render(self, pose, sharded: bool):
        from nerf_tex_b200.sharded import render_sharded
        params = self.params_host.to(self.dev, non_blocking=True)
        rays_o, rays_d, t, cone = self.sampler(None, pose)
        args = (rays_o[None], rays_d[None], t[None], params, cone[None])
        if sharded and self.world > 1:
            return render_sharded(self.r, *args, block=self.STRIPE, interleave=True, composite_bkgd=True, training=False)
        return self.r(*args, composite_bkgd=True, training=False)

    def workload_stats(self):
        """Ray / sample counts of pose 0 (what the frame actually contains)."""
        import torch
        rays_o, rays_d, t, cone = self.sampler(None, self.poses[0])
        hit = torch.isfinite(t[:, 0])
        ro, rd = rays_o[hit].contiguous(), rays_d[hit].contiguous()
        params = self.params_host.to(self.dev)
        n_s = n_h = 0
        for i in range(0, ro.shape[0], 16384):
            csr = self.inst.get_model_input_csr(ro[i:i + 16384], rd[i:i + 16384], params.repeat(ro[i:i + 16384].shape[0], 1),
                                                N_SAMPLES, STEP_SIZE, seed=0)
            n_s, n_h = n_s + csr["pts"].shape[0], n_h + csr["ray_index"].shape[0]
        return {"rays": H * W, "instances": self.n_inst, "proxy_hit_rays": int(ro.shape[0]), "shell_hit_rays": n_h,
                "in_shell_samples": n_s, "samples_per_shell_hit_ray": n_s / max(n_h, 1), "render_chunk": self.chunk}

    def instancer_chunk_ms(self):
        """The instancer alone (trace + march) on one chunk from the middle of the frame."""
        import torch
        rays_o, rays_d, t, cone = self.sampler(None, self.poses[0])
        hit = torch.isfinite(t[:, 0])
        mid, c = int(hit.sum()) // 2, self.chunk
        ro, rd = rays_o[hit][mid - c // 2:mid + c // 2].contiguous(), rays_d[hit][mid - c // 2:mid + c // 2].contiguous()
        pf = self.params_host.to(self.dev).repeat(ro.shape[0], 1)
        for _ in range(2):
            self.inst.get_model_input_csr(ro, rd, pf, N_SAMPLES, STEP_SIZE, seed=0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            self.inst.get_model_input_csr(ro, rd, pf, N_SAMPLES, STEP_SIZE, seed=0)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / 10

    def frame(self, i: int, sharded: bool = True):
        from nerf_tex_b200.image import finish_image
        out = self.render(self.poses[i % len(self.poses)], sharded)
        _, u8 = finish_image(out["color_pred"], out["alpha_pred"], H, W, 1, straight_alpha=True, want_u8=True)
        self.host_img.copy_(u8, non_blocking=True)
        return u8


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from nerf_tex_b200 import _lib
    from nerf_tex_b200.model import ParamNerf
    from nerf_tex_b200.renderer import InstanceRenderer
    from nerf_tex_b200.sharded import render_sharded, shard_index

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()
    STRIPE = DeviceFramePipeline.STRIPE

    ff = lambda n: {"module": "network.model.FourierFeatures", "n_freq_bands": n}
    model = ParamNerf(ff(10), ff(4), ff(4), n_parameters=list(NPAR), seed=1)["model"]
    rays_o, rays_d, t, params, cone = frame_rays(seed=1000)                  # ONE frame, the same on every rank
    n_rays = H * W

    # ---- the frame's instancer records (not part of the hot path), then this rank's share of them
    d_o, d_d = torch.as_tensor(rays_o[0]).to(dev), torch.as_tensor(rays_d[0]).to(dev)
    d_p = torch.as_tensor(params).to(dev).repeat_interleave(n_rays, dim=0)
    records = FrameRecords(dev, d_o, d_d, d_p)
    del d_p
    n_samples_total, n_hit = records.n_samples_total, records.n_hit
    my_rays = shard_index(n_rays, rank, world, STRIPE, True, dev) if world > 1 else torch.arange(n_rays, device=dev)
    chunks = records.chunks_for(my_rays)
    my_samples = int(sum(c["pts"].shape[0] for c in chunks))
    dev_in = [torch.as_tensor(x).to(dev) for x in (rays_o, rays_d, t, params, cone)]
    mk_renderer = lambda inst: InstanceRenderer(instancer_config=inst, model=model, n_samples=N_SAMPLES,
                                                render_chunk=RENDER_CHUNK, net_chunk=32768, step_size=STEP_SIZE,
                                                density_reweighting=True)
    resident = ChunkReplayInstancer(chunks, resident=True)
    renderer = mk_renderer(resident)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def one_step():
        """One frame: this rank's stripes through K1 / K2 chunk by chunk, then (N > 1) the final gather."""
        resident.reset()
        if world > 1:
            return render_sharded(renderer, *dev_in, block=STRIPE, interleave=True, composite_bkgd=False, training=False)
        return renderer(*dev_in, composite_bkgd=False, training=False)

    # ---- N > 1: the gathered frame must be the single-GPU frame, bit for bit (checked once, outside the timed region)
    checks = {}
    if world > 1:
        whole_inst = ChunkReplayInstancer(records.chunks_for(torch.arange(n_rays, device=dev)), resident=True)
        whole = mk_renderer(whole_inst)(*dev_in, composite_bkgd=False, training=False)
        parts = one_step()
        same = all(torch.equal(whole[k], parts[k]) for k in whole) and set(whole) == set(parts)
        flags = [None] * world
        dist.all_gather_object(flags, bool(same))
        checks["sharded_eq_single"] = all(flags)
        del whole, parts
    del records
    torch.cuda.empty_cache()

    # K1 launch timing hooks (CUDA events on the launching stream)
    k1_events = []
    orig_forward = model.forward_rows

    def timed_forward(**kw):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        out = orig_forward(**kw)
        e.record()
        k1_events.append((s, e, kw["n_rows"]))
        return out

    for _ in range(max(args.warmup, 3)):
        one_step()
    model.forward_rows = timed_forward
    sync_all()
    proc, cpath = clocks_start()
    time.sleep(0.3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    torch.cuda.nvtx.range_push("ntx_value_timed")        # ncu --nvtx --nvtx-include "ntx_value_timed/": the timed region only
    for _ in range(args.steps):
        out = one_step()
    torch.cuda.nvtx.range_pop()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1)
    model.forward_rows = orig_forward
    time.sleep(0.2)
    clocks = clocks_stop(proc, cpath)
    k1_ms = sum(s.elapsed_time(e) for s, e, _ in k1_events)
    k1_rows = sum(n for _, _, n in k1_events)
    k1_launches = len(k1_events)
    n_chunks = len(chunks)
    del chunks, resident, renderer

    # ---- N > 1, extra key: round 1's weak-scaling figure (every rank renders the whole frame on its own, no gather)
    ms_weak = 0.0
    if world > 1:
        whole_r = mk_renderer(whole_inst)

        def weak_step():
            whole_inst.reset()
            return whole_r(*dev_in, composite_bkgd=False, training=False)
        weak_step()
        sync_all()
        w_steps = max(1, min(args.steps, 3))
        e0.record()
        for _ in range(w_steps):
            weak_step()
        e1.record()
        sync_all()
        ms_weak = e0.elapsed_time(e1) / w_steps
        del whole_inst, whole_r
    torch.cuda.empty_cache()

    # ---- e2e arm: the device-instancer pipeline through the public API — pose + parameters in from the host, finished
    # frame out to pinned host memory, every frame, inside the timed region; sharded over the ranks for N > 1
    pipe = DeviceFramePipeline(dev, world)
    for i in range(3):
        pipe.frame(i)
    sync_all()
    e_steps = max(1, args.steps)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    torch.cuda.nvtx.range_push("ntx_e2e_timed")
    for i in range(e_steps):
        pipe.frame(i)
    torch.cuda.nvtx.range_pop()
    e3.record()
    sync_all()
    ms_e2e = e2.elapsed_time(e3)

    # ---- max over ranks
    stats = torch.tensor([ms, ms_e2e, k1_ms, ms_weak], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(k1_rows), float(my_samples)], dtype=torch.float64, device=dev)
    mx = torch.tensor([float(my_samples)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    ms, ms_e2e, k1_ms_max, ms_weak = [float(x) for x in stats.tolist()]
    k1_rows_all, samples_all = [float(x) for x in tot.tolist()]
    assert int(samples_all) == n_samples_total, (samples_all, n_samples_total)   # the shards tile the frame

    pipe_stats = pipe.workload_stats() if rank == 0 else {}
    h2d, d2h = pipe.h2d_bytes, pipe.d2h_bytes
    inst_ms = pipe.instancer_chunk_ms() if world == 1 else None
    del pipe
    torch.cuda.empty_cache()
    dp = None
    if world > 1:                      # every rank takes part (one gradient all-reduce per step)
        dp = train_step_dp_bench(dev, rank, world)
        checks.update(dp.pop("checks"))

    if rank == 0:
        pk = peaks()
        sec = ms / 1e3
        rays_per_s = n_rays * args.steps / sec
        samples_per_s = n_samples_total * args.steps / sec
        flops_per_row = 2.0 * MACS_PER_SAMPLE[NPAR]
        k1_tflops = (k1_rows / max(k1_launches, 1)) * flops_per_row / ((k1_ms / max(k1_launches, 1)) / 1e3) / 1e12
        traffic, traffic_src = k1_traffic()
        line = {
            "metric": "rays/s (800x800 carpet-shape instanced render; samples/s alongside)",
            "value": rays_per_s, "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": config_dict({"hit_rays_per_frame": n_hit, "in_shell_samples_per_frame": n_samples_total,
                                   "parallelism": (f"one frame over {world} GPUs: round-robin stripes of {STRIPE} rays, no collective "
                                                   f"but the final all-gather of the per-ray outputs (inside the timed region)")
                                   if world > 1 else "one GPU",
                                   "load_balance": {"max_rank_samples": float(mx), "mean_rank_samples": samples_all / world}}),
            "samples_per_sec": samples_per_s,
            "roofline": {"kernel": "mlp_forward_kernel (K1 fused encoding + MLP, tcgen05)", "bound": "tensor",
                         "achieved": k1_tflops, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                         "frac": k1_tflops / pk["bf16_tflops_sustained"], "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": f"{pk['source']} bf16_tflops_sustained (kernel timed inside a long step)",
                         "frac_of_burst_peak": k1_tflops / pk["bf16_tflops"],
                         "flops_per_sample": flops_per_row, "rows_per_launch": k1_rows / max(k1_launches, 1),
                         "launch_ms": k1_ms / max(k1_launches, 1), "k1_share_of_step": k1_ms_max / ms},
            "e2e": {"value": n_rays * e_steps / (ms_e2e / 1e3), "unit": "rays/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / e_steps,
                    "workload": pipe_stats,
                    "api": "Proxy ray sampler -> InstanceRenderer(Instancer(1500-patch procedural carpet)) -> finish_image "
                           "[render_sharded for N > 1]: camera pose + parameter vector in from the host, finished 8-bit RGBA "
                           "frame out to pinned host memory, every frame"},
            "gpu_launches": int(2 * n_chunks * args.steps),
            "clocks": clocks,
        }
        if world > 1:
            line["checks"] = checks
            line["train_step_dp"] = dp
            line["weak_scaling"] = {"value": world * n_rays / (ms_weak / 1e3), "unit": "rays/s", "ms_per_step": ms_weak,
                                    "what": "every rank renders the whole frame on its own (round 1's N > 1 figure)"}
        if world == 1 and not args.no_cpu_baseline:
            line["e2e"]["instancer_ms_per_131072_ray_chunk"] = inst_ms
            line["e2e_host_records"] = host_records_e2e(dev, model, rays_o, rays_d, t, params, cone, min(args.steps, 3))
            line["hbm_kernels"] = hbm_kernels_bench(dev, pk)
            line["train_step"] = train_step_bench(dev)
            line["c1_forward"] = c1_forward_bench(dev)
            line["c5_plush_frame"] = c5_plush_bench(dev)
            line["cpu_baseline"] = cpu_baseline_run(steps=3, warmup=1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def host_records_e2e(dev, model, rays_o, rays_d, t, params, cone, steps):
    """Round 1's end-to-end arm, kept for continuity: the same frame with a HOST-side instancer — rays and per-chunk
    instancer records copied from pinned host memory inside the timed region (3.9 GB per frame), colour / alpha read back."""
    import torch
    from nerf_tex_b200.renderer import InstanceRenderer
    n_rays = H * W
    d_o, d_d = torch.as_tensor(rays_o[0]).to(dev), torch.as_tensor(rays_d[0]).to(dev)
    d_p = torch.as_tensor(params).to(dev).repeat_interleave(n_rays, dim=0)
    records = FrameRecords(dev, d_o, d_d, d_p)
    chunks = records.chunks_for(torch.arange(n_rays, device=dev))
    del records, d_p
    # `t` feeds only the blur_idx scaling (renderer.py:260-263) and `instance_id` only false-colour rendering
    # (:306-307); neither is enabled in the carpet config, so the host does not send them
    needed = {"ray_ptr", "ray_index", "pts", "dirs", "params", "dists", "alpha_weight", "color_last", "alpha_last"}
    host = ChunkReplayInstancer(chunks, resident=False, fields=needed)
    del chunks
    torch.cuda.empty_cache()
    renderer_h = InstanceRenderer(instancer_config=host, model=model, n_samples=N_SAMPLES, render_chunk=RENDER_CHUNK,
                                  net_chunk=32768, step_size=STEP_SIZE, density_reweighting=True)
    host_in = [torch.as_tensor(x).pin_memory() for x in (rays_o, rays_d, t, params, cone)]
    res_c = torch.empty((1, n_rays, 3), dtype=torch.float32).pin_memory()
    res_a = torch.empty((1, n_rays), dtype=torch.float32).pin_memory()

    def step():
        host.reset()
        ins = [x.to(dev, non_blocking=True) for x in host_in]
        o = renderer_h(*ins, composite_bkgd=False, training=False)
        res_c.copy_(o["color_pred"], non_blocking=True)
        res_a.copy_(o["alpha_pred"], non_blocking=True)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    host.h2d_bytes = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": n_rays / (ms / 1e3), "unit": "rays/s", "ms_per_step": ms,
            "h2d_bytes_per_step": int(host.h2d_bytes // steps + sum(x.numel() * 4 for x in host_in)),
            "d2h_bytes_per_step": int(res_c.numel() * 4 + res_a.numel() * 4),
            "api": "InstanceRenderer.__call__ with pinned host rays + host instancer records (synthetic, as `value`)"}


def k1_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one K1 chunk launch from the newest committed ncu capture
    (profiles/*_k1_traffic.json, written by tools/ncu_traffic.py from `ncu --set full` of this very command)."""
    cands = sorted((ROOT / "profiles").glob("*_k1_traffic.json"))
    if cands:
        d = json.loads(cands[-1].read_text())
        return d["dram_bytes_per_launch"], f"profiles/{cands[-1].name}: {d['how']}"
    return K1_NCU_DRAM_BYTES_PER_LAUNCH, "ncu --set full capture of one chunk launch (profiles/r01_v6_k1_full_raw.csv)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
