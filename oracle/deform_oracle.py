"""TEST INFRASTRUCTURE -- CPU restatement of the reference's deformation network.

Plain functional torch (any device / dtype, runs in seconds on CPU at test sizes) following
scene/deformation.py of the reference line by line; it is the checker for the CUDA path and the
timed CPU baseline of bench.py. It is pinned against the reference module itself (imported by path
in the dev container) and against the golden vectors that import produced
(tests/golden/deform_golden.pt, script tests/golden/make_deform_golden.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
"""
import numpy as np
import torch
import torch.nn.functional as F

HEADS = ("pos", "scales", "rotations", "opacity", "rgb")


def get_temporal_embed(weight, t, n_rows):
    """deformation.py:53-67 -- returns the single [dim] row (the reference repeats it N times)."""
    dim = weight.shape[1]
    emb_resized = F.interpolate(weight[None, None, ...], size=(n_rows, dim), mode="bilinear", align_corners=True)
    col = torch.arange(dim, device=weight.device, dtype=weight.dtype).unsqueeze(-1) / (dim - 1)
    grid = torch.cat([col, torch.ones(dim, 1, device=weight.device, dtype=weight.dtype) * t], dim=-1)[None, None, ...]
    grid = (grid - 0.5) * 2
    emb = F.grid_sample(emb_resized, grid, align_corners=True, mode="bilinear", padding_mode="reflection")
    return emb.reshape(dim)


def int_lininterp(t, init_val, final_val, until):
    """deformation.py:69-70"""
    return int(init_val + (final_val - init_val) * min(max(t, 0), until) / until)


def _seq(sd, prefix, x, n_linear, first_relu):
    """nn.Sequential of Linear layers at even (feature_out) or odd (heads) indices with ReLU between."""
    idx = 1 if first_relu else 0
    for i in range(n_linear):
        if first_relu or i > 0:
            x = torch.relu(x)
        x = F.linear(x, sd[f"{prefix}.{idx}.weight"], sd[f"{prefix}.{idx}.bias"])
        idx += 2
    return x


def query_time(sd, args, stage, t, emb, it, num_down_emb, use_coarse, depth, max_embeddings, c2f_iter):
    """deformation.py:72-88"""
    if use_coarse:
        n = num_down_emb
    elif args.no_c2f_temporal_embedding:
        n = max_embeddings
    else:
        n = int_lininterp(it, num_down_emb, max_embeddings, c2f_iter)
    tau = get_temporal_embed(sd["weight"], t, n)
    h = torch.cat([tau.unsqueeze(0).expand(emb.shape[0], -1), emb], dim=-1)
    return _seq(sd, f"feature_out_{stage}", h, depth, first_relu=False)


def deform(sd, args, stage, hidden, pts, scales, rotations, opacity, sh, scale, scale_c, scale_o, coef_s):
    """deformation.py:90-106"""
    pts = pts + _seq(sd, f"pos_deform_{stage}", hidden, 2, True) * scale
    if not args.no_ds:
        scales = scales + _seq(sd, f"scales_deform_{stage}", hidden, 2, True) * scale * coef_s
    if not args.no_dr:
        rotations = rotations + _seq(sd, f"rotations_deform_{stage}", hidden, 2, True) * scale
    if not args.no_do:
        opacity = opacity + _seq(sd, f"opacity_deform_{stage}", hidden, 2, True) * scale * scale_o
    if not args.no_dc:
        sh = sh + _seq(sd, f"rgb_deform_{stage}", hidden, 2, True).view(-1, 16, 3) * scale_c
    return pts, scales, rotations, opacity, sh


def forward(sd, args, depth, max_embeddings, point, scales, rotations, opacity, time_emb, cam_no, embedding, sh, it,
            num_down_emb_c=30, num_down_emb_f=30):
    """deformation.py:108-141. `sd` maps the reference's parameter names to tensors; time_emb is
    modified in place like the reference does."""
    pts, scales, rotations, opacity = point[:, :3], scales[:, :3], rotations[:, :4], opacity[:, :1]
    orig = (pts, scales, rotations, opacity, sh)
    if cam_no is None:
        off = sd["offsets"]
        offset = torch.masked_select(off, off.ne(0)).mean()
        offset = torch.where(torch.isnan(offset), torch.zeros_like(offset), offset)
    else:
        offset = sd["offsets"][cam_no]
    time_emb += offset
    t = time_emb[0, 0]
    use_anneal = args.use_anneal
    coef = 1 if not use_anneal else float(np.clip(it / 1000, 0, 1))
    coef_c = 1 if not use_anneal else float(np.clip((it - args.deform_from_iter) / 1000, 0, 1))
    coef_o = coef_c
    coef_s = coef_c
    c2f = args.c2f_temporal_iter
    if args.no_coarse_deform:
        sub = orig
    else:
        hidden = query_time(sd, args, "c", t, embedding, it, num_down_emb_c, args.use_coarse_temporal_embedding, depth,
                            max_embeddings, c2f)
        sub = deform(sd, args, "c", hidden, pts, scales, rotations, opacity, sh, coef, coef_c, coef_o, coef_s)
    if args.no_fine_deform:
        out = sub
    else:
        hidden = query_time(sd, args, "f", t, embedding, it, num_down_emb_f, False, depth, max_embeddings, c2f)
        out = deform(sd, args, "f", hidden, *sub, coef, coef_c, coef_o, coef_s)
    return (*out, (sub, orig))
