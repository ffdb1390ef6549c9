"""Out-of-bounds-write check of our own (compute-sanitizer is closed on this pool, profiles/r02_sanitizer_unavailable.txt).

Every device buffer the host wrappers allocate for a kernel (outputs, workspaces, saved operands) is replaced, for the
duration of a call, by a slice of a larger allocation with 4 KiB guard bands poisoned with 0xA5 on both sides; after the
call (and a synchronize) every band must still hold the poison.  Shapes are deliberately not multiples of the kernels'
tile sizes (128-row MLP tiles, 32-lane rounds, 16-byte vector accesses), where an unguarded tail store would land.
The results themselves are checked by the parity tests; this file only looks at what lies next to them."""

import contextlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096
POISON = 0xA5


class Guards:
    def __init__(self):
        self.bands = []
        self._orig = {}

    def _alloc(self, shape, dtype, device, fill=None):
        if isinstance(shape, int):
            shape = (shape,)
        shape = tuple(int(s) for s in shape)
        n = int(np.prod(shape)) if len(shape) else 1
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        body = (nbytes + 255) // 256 * 256
        raw = self._orig["empty"](GUARD + body + GUARD, dtype=torch.uint8, device=device)
        raw.fill_(POISON)
        self.bands.append((raw, nbytes))
        t = raw[GUARD:GUARD + nbytes].view(dtype).view(shape)
        if fill is not None:
            t.fill_(fill)
        return t

    def install(self):
        g = self

        def is_cuda(device):
            return device is not None and torch.device(device).type == "cuda"

        def empty(*size, dtype=torch.float32, device=None, **kw):
            if not is_cuda(device) or kw:
                return g._orig["empty"](*size, dtype=dtype, device=device, **kw)
            return g._alloc(size[0] if len(size) == 1 and not isinstance(size[0], int) else size, dtype, device)

        def zeros(*size, dtype=torch.float32, device=None, **kw):
            if not is_cuda(device) or kw:
                return g._orig["zeros"](*size, dtype=dtype, device=device, **kw)
            return g._alloc(size[0] if len(size) == 1 and not isinstance(size[0], int) else size, dtype, device, fill=0)

        def empty_like(t, **kw):
            if not t.is_cuda or kw:
                return g._orig["empty_like"](t, **kw)
            return g._alloc(tuple(t.shape), t.dtype, t.device)

        def zeros_like(t, **kw):
            if not t.is_cuda or kw:
                return g._orig["zeros_like"](t, **kw)
            return g._alloc(tuple(t.shape), t.dtype, t.device, fill=0)

        for name, fn in (("empty", empty), ("zeros", zeros), ("empty_like", empty_like), ("zeros_like", zeros_like)):
            self._orig[name] = getattr(torch, name)
            setattr(torch, name, fn)

    def remove(self):
        for name, fn in self._orig.items():
            setattr(torch, name, fn)

    def check(self, tag):
        torch.cuda.synchronize()
        assert self.bands, f"{tag}: no guarded allocation was made (the wrappers changed?)"
        for raw, nbytes in self.bands:
            front = raw[:GUARD]
            back = raw[GUARD + nbytes:]
            assert bool((front == POISON).all()), f"{tag}: write BEFORE a {nbytes}-byte buffer"
            assert bool((back == POISON).all()), f"{tag}: write PAST the end of a {nbytes}-byte buffer"


@contextlib.contextmanager
def guarded(tag):
    g = Guards()
    g.install()
    try:
        yield g
    finally:
        g.remove()
    g.check(tag)


def _rays(R, p, seed=0):
    gen = torch.Generator(device="cuda").manual_seed(seed)
    ro = torch.tensor([[0., 0.3, 5.]], device="cuda").repeat(R, 1)
    rd = torch.rand((R, 3), device="cuda", generator=gen) * 2 - 1 - ro
    rd = rd / rd.norm(dim=-1, keepdim=True)
    t = torch.stack([3 + torch.rand(R, device="cuda", generator=gen), 6 + torch.rand(R, device="cuda", generator=gen)], -1)
    return ro, rd, t, torch.rand((R, p), device="cuda", generator=gen), torch.rand((R, 1), device="cuda", generator=gen) * .01, gen


@pytest.mark.parametrize("R,S,N", [(1, 3, 1), (37, 64, 128), (131, 63, 77), (1000, 64, 128), (5, 255, 40)])
def test_sampling_kernels_stay_in_bounds(R, S, N):
    from nerf_tex_b200 import ops
    ro, rd, t, _, _, gen = _rays(R, 5)
    z_rand = torch.rand((R, S), device="cuda", generator=gen)
    w = torch.rand((R, S), device="cuda", generator=gen) ** 6
    u = torch.rand((R, N), device="cuda", generator=gen)
    with guarded("stratified_z"):
        z = ops.stratified_z(t, S, True, z_rand)
    for det in (True, False):
        with guarded(f"hierarchical_z det={det}"):
            ops.hierarchical_z(z, w, N, det, None if det else u)
        with guarded(f"sample_pdf det={det}"):
            ops.sample_pdf(z[:, :-1].contiguous(), w[:, :S - 2].contiguous(), N, det, None if det else u)


@pytest.mark.parametrize("R,S", [(1, 2), (7, 33), (300, 256), (33, 1023), (4097, 64)])
def test_compositing_kernels_stay_in_bounds(R, S):
    from nerf_tex_b200 import ops
    ro, rd, t, _, _, gen = _rays(R, 5)
    rgbs = torch.randn((R, S, 4), device="cuda", generator=gen)
    z = torch.sort(torch.rand((R, S), device="cuda", generator=gen) * 3 + 2, -1).values
    with guarded("composite_forward"):
        cmap, amap, w, _ = ops.composite_forward(rgbs, z, rd, want_weights=True, composite_bkgd=True)
    with guarded("composite_backward"):
        ops.composite_backward(rgbs, z, rd, torch.randn_like(cmap), torch.randn_like(amap))
    if S >= 3:
        with guarded("composite_edges_forward"):
            ops.composite_edges_forward(rgbs[:, :S - 1].contiguous(), z, rd, want_weights=True)


def test_instanced_compositor_stays_in_bounds():
    from nerf_tex_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(3)
    n_rays = 1234
    counts = torch.randint(0, 70, (777,), device="cuda", generator=gen)
    counts[::50] = 0
    ptr = torch.zeros(778, dtype=torch.int64, device="cuda")
    ptr[1:] = torch.cumsum(counts, 0)
    T = int(ptr[-1])
    ray_index = torch.randperm(n_rays, device="cuda", generator=gen)[:777].sort().values.to(torch.int32)
    with guarded("composite_instanced_forward"):
        ops.composite_instanced_forward(torch.randn((T, 4), device="cuda", generator=gen),
                                        torch.rand(T, device="cuda", generator=gen) * 0.002,
                                        torch.rand(T, device="cuda", generator=gen), ptr,
                                        torch.rand((777, 3), device="cuda", generator=gen),
                                        (torch.rand(777, device="cuda", generator=gen) < .5).float(), ray_index, n_rays,
                                        patch_scale=0.09, composite_bkgd=True)


@pytest.mark.parametrize("n,npar", [(1, (1, 4)), (129, (2, 3)), (255, (1, 6)), (1000, (0, 0)), (5000, (1, 4))])
def test_fused_mlp_stays_in_bounds(n, npar):
    from nerf_tex_b200.model import FusedNerfTexMLP
    m = FusedNerfTexMLP(n_geo=npar[0], n_app=npar[1], color_depth=0 if npar == (0, 0) else 1, seed=1)
    gen = torch.Generator(device="cuda").manual_seed(n)
    pos = torch.rand((n, 3), device="cuda", generator=gen) * 2 - 1
    d = torch.nn.functional.normalize(torch.randn((n, 3), device="cuda", generator=gen), dim=-1)
    prm = torch.rand((n, sum(npar)), device="cuda", generator=gen) if sum(npar) else None
    m.ensure_packed()
    with guarded("mlp_forward"):
        m.forward_rows(n_rows=n, pos=pos, dir=d, params=prm, rows_per_group=1)


@pytest.mark.parametrize("R,S", [(3, 64), (70, 100), (200, 256)])
def test_training_pass_stays_in_bounds(R, S):
    """Forward with saved tile images, compositing backward, fused dgrad chain, wgrad and heads kernels, Adam."""
    from nerf_tex_b200.loss import AlphaLoss
    from nerf_tex_b200.model import FusedNerfTexMLP
    from nerf_tex_b200.renderer import Renderer
    from nerf_tex_b200.train import Trainer
    m = FusedNerfTexMLP(n_geo=2, n_app=3, seed=2)
    ro, rd, t, prm, cone, gen = _rays(R, 5, seed=R)
    r = Renderer(model=m, n_samples=S, perturb=True, blur_idx=0, raw_noise_std=0.1, seed=0)
    tr = Trainer({"model": m}, r, AlphaLoss("network.loss.smape", "network.loss.mse"))
    data = {"rays_o": ro[None], "rays_d": rd[None], "t": t[None], "parameters": prm[:1], "cone_scale": cone[None],
            "color": torch.rand((1, R, 3), device="cuda", generator=gen), "alpha": torch.rand((1, R), device="cuda", generator=gen)}
    tr.train_step(data)                                     # packs weights, builds the schedule, fills the pools
    with guarded("train_step"):
        tr.train_step(data)


def test_chunk_call_instancer_rays_and_image_stay_in_bounds():
    from nerf_tex_b200 import ops
    from nerf_tex_b200.image import finish_image
    from nerf_tex_b200.instancer import Instancer
    from nerf_tex_b200.model import FusedNerfTexMLP
    R = 333
    ro, rd, t, prm, cone, gen = _rays(R, 5, seed=4)
    coarse, fine = FusedNerfTexMLP(1, 4, seed=1), FusedNerfTexMLP(1, 4, seed=2)
    cfg = ops.render_config(64, 128, True, False, True, None, (1., 1., 1.))
    z_rand = torch.rand((R, 64), device="cuda", generator=gen)
    coarse.ensure_packed(), fine.ensure_packed()
    with guarded("render_rays"):
        ops.render_rays(coarse, fine, cfg, ro, rd, t, prm, cone, z_rand=z_rand)
    T = np.tile(np.eye(4), (40, 1, 1))
    rng = np.random.default_rng(0)
    T[:, :3, :3] *= 0.2
    T[:, :3, 3] = rng.uniform(-1, 1, (40, 3))
    inst = Instancer(b_0=[-1, -1, -1], b_1=[1, 1, 1], textures=["", "", "light"], transformations=T, patch_scale=0.2,
                     instance_sampling_method="nearest")
    with guarded("instancer"):
        inst.get_model_input_csr(ro, rd, prm, 257, 0.01, seed=1)
    c2w = np.eye(4, dtype=np.float32)
    c2w[2, 3] = 4.0
    with guarded("rays_from_camera"):
        ops.rays_from_camera(None, 37, 53, 60.0, c2w, normalize=True, aabb=([-1, -1, -1], [1, 1, 1]))
    with guarded("image_finish"):
        finish_image(torch.rand((37 * 53, 3), device="cuda"), torch.rand(37 * 53, device="cuda"), 37, 53, 2, want_u8=True)
