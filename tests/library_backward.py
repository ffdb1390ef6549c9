"""Layer-wise library-GEMM backward (torch.mm -> cuBLAS, bf16 in / fp32 accumulate) on the operands the fused forward
saved: test infrastructure that cross-checks the hand-written backward kernels (nerf_tex_b200/autograd.py::
mlp_backward_native) on identical inputs.  Not part of the product path."""

import torch

from nerf_tex_b200.autograd import ENC_COLS, untile


def mlp_backward(model, saved, d_rgbs: torch.Tensor) -> torch.Tensor:
    """d loss / d master given d loss / d rgbs[N,4]  (model.py:104-123 in reverse).

    All GEMMs take bf16 operands with fp32 accumulation; weight gradients come out in fp32.  (An
    fp32 / TF32 gradient chain was measured to give the same agreement with the float64 reference —
    the residual is the bf16 forward itself — so the cheaper bf16 chain is kept.)"""
    enc_pos, enc_dir, acts_img, masks, acts, n_rows = saved
    enc_pos, enc_dir = untile(enc_pos, n_rows, ENC_COLS), untile(enc_dir, n_rows, ENC_COLS)
    if acts is None:
        acts = [untile(acts_img[st], n_rows) for st in range(acts_img.shape[0])]
    sched = model.schedule()
    master = model.master.detach()
    grad = torch.zeros_like(master)
    bf = torch.bfloat16

    def wview(layer):
        wo, bo, fi, fo = model.layers[layer]
        return master[wo:wo + fi * fo].view(fi, fo), grad[wo:wo + fi * fo].view(fi, fo), grad[bo:bo + fo]

    d_rgb = d_rgbs[:, :3].contiguous()
    d_sigma = d_rgbs[:, 3:4].contiguous()

    # colour head: rgb = h_half @ W_rgb + b   (model.py:123)
    last = sched.n_steps - 1
    h_half = acts[last][:, :128]
    W, gW, gb = wview(sched.rgb_layer)
    gW.copy_(torch.mm(h_half.t(), d_rgb.to(bf), out_dtype=torch.float32))
    gb.copy_(d_rgb.sum(0))
    d_out = torch.mm(d_rgb.to(bf), W.t().to(bf))                        # [N,128] bf16

    # density head: sigma = h_trunk @ w_alpha + b   (model.py:111)
    Wa, gWa, gba = wview(sched.alpha_layer)
    h_trunk = acts[sched.depth - 1]
    gWa.copy_(torch.mm(h_trunk.t(), d_sigma.to(bf), out_dtype=torch.float32))
    gba.copy_(d_sigma.sum(0))

    for st in range(last, -1, -1):
        s = sched.steps[st]
        n_out = s["n_out"]
        if st == sched.depth - 1:                                        # alpha head branches off the trunk output
            d_out = d_out + torch.mm(d_sigma.to(bf), Wa.t().to(bf))
        act = acts[st][:, :n_out]
        dY = torch.where(act > 0, d_out, torch.zeros_like(d_out)) if s["relu"] else d_out
        dYb = dY
        W, gW, gb = wview(s["layer"])
        gb.copy_(dY.sum(0, dtype=torch.float32))
        r0 = s["main_row0"]
        if st > 0:                                                       # main operand = previous step's stored output
            x = acts[st - 1]
            gW[r0:r0 + 256].copy_(torch.mm(x.t(), dYb, out_dtype=torch.float32))
        if s["enc_data"]:                                                # encoded inputs (pos at trunk0 / skip, dir at colour)
            e = enc_pos if s["phase"] == 0 else enc_dir
            cols, rows = sched.enc_maps[s["phase"]]
            gE = torch.mm(e.t(), dYb, out_dtype=torch.float32)           # [96, n_out] in kernel column order
            gW[rows] = gE[cols]
        if st > 0:
            d_out = torch.mm(dY, W[r0:r0 + 256].t().to(bf))              # dgrad into the previous step's output
    return grad
