"""Drop-in for the reference's ``scene.deformation.deform_network`` (scene/deformation.py:15-148).

Same constructor, same ``forward`` signature and return structure, same parameter names (so
``deformation.pth`` state dicts load both ways), same in-place ``time_emb += offset`` side effect.
The arithmetic runs in libed3dgs_b200.so: temporal-embedding kernel + fused tensor-core MLP
(3xTF32, fp32-grade) forward and backward, behind one autograd node.
"""
import ctypes
import weakref

import numpy as np
import torch
import torch.nn as nn

from . import _deform_abi as abi
from . import _lib

# Multi-GPU overlap (dp.OverlapSync): when a side stream is set, the kernels of the network's backward are issued on
# it, so that the all-reduce of the Gaussian-parameter gradients (which are this node's *upstream* gradients and hence
# final before it starts) can run on NCCL's stream at the same time. Everything the side stream still reads or writes
# is parked in `keep` until dp.OverlapSync has made the main stream wait for it.
#
# Restriction, and what guards it. The gradients this node returns for its leaves (embedding, temporal table, camera
# offsets, Linear parameters) are still being written by the side stream when backward() returns, while autograd
# believes they belong to the main stream. That is only sound while nobody touches them on the main stream before
# dp.OverlapSync.join(): one deformation backward per step, every leaf with .grad unset (AccumulateGrad then adopts
# the tensor without reading it), and no other term of the loss feeding the same leaves. The first two are checked in
# backward(): a second backward before the join, or a leaf that already holds a .grad, makes the main stream wait for
# the side stream before the gradients are handed back (`serialised` counts those). The third cannot be seen from
# inside the node (autograd sums the contributions of several nodes in its input buffer, on the main stream): a caller
# whose loss has such terms -- the reference's embedding regulariser on _embedding, train.py:229-233, or its temporal
# smoothness term on the table, train.py:236-242 -- must say so with shared_leaves=True, which also serialises.
_overlap = {"stream": None, "sm_reserve": 0, "clone": True, "keep": [], "shared_leaves": False, "serialised": 0}


def set_backward_stream(stream, sm_reserve=0, clone=True, shared_leaves=False):
    """stream = None switches the overlap off. sm_reserve: SMs the persistent backward kernels leave to the exchange.
    clone: the kernels read copies of the pass-through gradients (needed when the caller reduces them in place).
    shared_leaves: other terms of the loss also reach the embedding / table / offsets / network parameters in the same
    backward pass; the main stream then waits for the side stream before this node's gradients leave it."""
    _overlap["stream"] = stream
    _overlap["sm_reserve"] = int(sm_reserve) if stream is not None else 0
    _overlap["clone"] = bool(clone)
    _overlap["shared_leaves"] = bool(shared_leaves) if stream is not None else False
    _overlap["serialised"] = 0
    _overlap["keep"].clear()


def _must_serialise(leaf_refs):
    """True when the gradients of an overlapped backward may be read on the main stream before the join."""
    if _overlap["shared_leaves"] or _overlap["keep"]:
        return True
    for r in leaf_refs:
        leaf = r()
        if leaf is not None and leaf.grad is not None:
            return True
    return False


_HEAD_NAMES = ("pos", "scales", "rotations", "opacity", "rgb")
# deform_network.precision -> ed3dgs_deform_weights::precision (include/ed3dgs_deform.h)
_PRECISIONS = {"fp32": 0, "3xtf32": 0, "tf32": 1}


def _stage_tensors(feature_out, heads):
    """Flat list of the Linear parameters of one stage in a fixed order + its structure."""
    feat = [m for m in feature_out if isinstance(m, nn.Linear)]
    tensors = []
    for m in feat:
        tensors += [m.weight, m.bias]
    for seq in heads:
        tensors += [seq[1].weight, seq[1].bias, seq[3].weight, seq[3].bias]
    return tensors, len(feat)


def _weights_struct(tensors, depth, W, temb, gemb, enabled, precision=0):
    s = abi.DeformWeights()
    s.W, s.depth, s.temb_dim, s.gemb_dim = W, depth, temb, gemb
    s.precision = precision
    for l in range(depth):
        s.feat_w[l] = tensors[2 * l].data_ptr()
        s.feat_b[l] = tensors[2 * l + 1].data_ptr()
    o = 2 * depth
    for h in range(5):
        if enabled[h]:
            s.head_w1[h] = tensors[o + 4 * h].data_ptr()
            s.head_b1[h] = tensors[o + 4 * h + 1].data_ptr()
            s.head_w2[h] = tensors[o + 4 * h + 2].data_ptr()
            s.head_b2[h] = tensors[o + 4 * h + 3].data_ptr()
    return s


def _grads_struct(grads, depth, enabled):
    s = abi.DeformGrads()
    for l in range(depth):
        s.feat_w[l] = grads[2 * l].data_ptr()
        s.feat_b[l] = grads[2 * l + 1].data_ptr()
    o = 2 * depth
    for h in range(5):
        if enabled[h]:
            s.head_w1[h] = grads[o + 4 * h].data_ptr()
            s.head_b1[h] = grads[o + 4 * h + 1].data_ptr()
            s.head_w2[h] = grads[o + 4 * h + 2].data_ptr()
            s.head_b2[h] = grads[o + 4 * h + 3].data_ptr()
    return s


def temporal_embed(weight, n_rows, t_dev):
    """tau[dim] at device-scalar time t (deformation.py:53-67, without the [N,dim] repeat)."""
    max_emb, dim = weight.shape
    tau = torch.empty(dim, dtype=torch.float32, device=weight.device)
    _lib.check(_lib.lib().ed3dgs_temporal_embed_forward(weight.data_ptr(), max_emb, dim, int(n_rows),
                                                        t_dev.data_ptr(), tau.data_ptr(), _lib.stream_handle()))
    return tau


class _DeformFunction(torch.autograd.Function):
    """inputs: cfg (dict), t (1-element CUDA tensor), the per-camera time offsets, temporal table, embedding, 5
    per-Gaussian inputs, then the coarse and fine Linear parameters. outputs: 5 coarse (sub) + 5 final tensors.
    `offsets` takes no part in the forward arithmetic (the caller has already added the camera's offset to t,
    deformation.py:122-127 of the reference); it is an input so that its gradient -- d_t routed to row cam_no, or spread
    over the non-zero rows when cam_no is None -- leaves this node finished, on whichever stream the backward runs."""

    @staticmethod
    def forward(ctx, cfg, t, offsets, table, emb, pts, scales, rots, opac, sh, *params):
        dev = emb.device
        N = emb.shape[0]
        lib = _lib.lib()
        nc = cfg["n_params_c"]
        pc, pf = list(params[:nc]), list(params[nc:])
        pc = [p.detach().contiguous() for p in pc]
        pf = [p.detach().contiguous() for p in pf]
        W, depth, temb, gemb = cfg["W"], cfg["depth"], cfg["temb"], cfg["gemb"]
        table_c = table.detach().contiguous()
        t_c = t.detach().reshape(1).float().contiguous()
        emb_c = emb.detach().contiguous().float()
        ins = [x.detach().contiguous().float() for x in (pts, scales, rots, opac, sh.reshape(N, 48))]
        wc = _weights_struct(pc, depth, W, temb, gemb, cfg["enabled"], cfg["precision"]) if cfg["coarse"] else None
        wf = _weights_struct(pf, depth, W, temb, gemb, cfg["enabled"], cfg["precision"]) if cfg["fine"] else None
        with torch.cuda.device(dev):
            tau_c = temporal_embed(table_c, cfg["rows_c"], t_c) if cfg["coarse"] else torch.zeros(temb, device=dev)
            tau_f = temporal_embed(table_c, cfg["rows_f"], t_c) if cfg["fine"] else torch.zeros(temb, device=dev)
            nbytes = lib.ed3dgs_deform_workspace_bytes(W, depth, N)
            work = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            subs = [torch.empty_like(x) for x in ins]
            outs = [torch.empty_like(x) for x in ins]
            in_arr = abi.PtrArr5(*[x.data_ptr() for x in ins])
            sub_arr = abi.PtrArr5(*[x.data_ptr() for x in subs])
            out_arr = abi.PtrArr5(*[x.data_ptr() for x in outs])
            coef = abi.FloatArr5(*cfg["coef"])
            if N > 0:
                _lib.check(lib.ed3dgs_deform_forward(
                    ctypes.byref(wc) if wc else None, ctypes.byref(wf) if wf else None, N, emb_c.data_ptr(),
                    tau_c.data_ptr(), tau_f.data_ptr(), ctypes.byref(in_arr), ctypes.byref(sub_arr),
                    ctypes.byref(out_arr), ctypes.byref(coef), work.data_ptr(), nbytes, _lib.stream_handle()))
        ctx.cfg = cfg
        # weak references only: the node must not keep its own inputs alive (see _must_serialise)
        ctx.leaf_refs = [weakref.ref(x) for x in (offsets, table, emb, *params)
                         if isinstance(x, torch.Tensor) and x.requires_grad and x.is_leaf]
        ctx.set_materialize_grads(False)   # unused outputs arrive as None instead of N-sized zero tensors
        ctx.save_for_backward(t_c, table_c, emb_c, tau_c, tau_f, work, offsets.detach(), *pc, *pf)
        subs[4] = subs[4].view(N, 16, 3)
        outs[4] = outs[4].view(N, 16, 3)
        return (*subs, *outs)

    @staticmethod
    def backward(ctx, *gouts):
        cfg = ctx.cfg
        t_c, table_c, emb_c, tau_c, tau_f, work, offsets = ctx.saved_tensors[:7]
        params = list(ctx.saved_tensors[7:])
        nc = cfg["n_params_c"]
        pc, pf = params[:nc], params[nc:]
        dev = emb_c.device
        N = emb_c.shape[0]
        lib = _lib.lib()
        W, depth, temb, gemb = cfg["W"], cfg["depth"], cfg["temb"], cfg["gemb"]
        dims = abi.HEAD_DIMS

        def prep(g, h):
            if g is None:
                return None
            return g.reshape(N, dims[h]).contiguous().float()

        d_sub = [prep(gouts[h], h) for h in range(5)]
        d_out = [prep(gouts[5 + h], h) for h in range(5)]
        # ONE zero fill for everything the kernels accumulate into: the parameter gradients of both stages, the two
        # temporal-row gradients, the table gradient and d_t are slices of one buffer (every slice starts on a
        # 256-byte boundary). With .grad unset (zero_grad(set_to_none=True), train.py:448) autograd adopts these
        # slices as the parameters' .grad without a copy.
        pad = lambda n: (n + 63) // 64 * 64  # noqa: E731
        sizes = [p.numel() for p in pc] + [p.numel() for p in pf] + [temb, temb, table_c.numel(), 1, offsets.numel()]
        flat = torch.zeros(sum(pad(n) for n in sizes), dtype=torch.float32, device=dev)
        offs, o = [], 0
        for n in sizes:
            offs.append(o)
            o += pad(n)
        gc = [flat[offs[i]:offs[i] + p.numel()].view(p.shape) for i, p in enumerate(pc)]
        gf = [flat[offs[len(pc) + i]:offs[len(pc) + i] + p.numel()].view(p.shape) for i, p in enumerate(pf)]
        k = len(pc) + len(pf)
        d_tau_c = flat[offs[k]:offs[k] + temb]
        d_tau_f = flat[offs[k + 1]:offs[k + 1] + temb]
        d_table = flat[offs[k + 2]:offs[k + 2] + table_c.numel()].view(table_c.shape)
        d_t = flat[offs[k + 3]:offs[k + 3] + 1]
        d_offsets = flat[offs[k + 4]:offs[k + 4] + offsets.numel()].view(offsets.shape)
        # the first stage that runs overwrites d_emb (accumulate flag), so it needs no zero fill
        d_emb = (torch.empty if (N > 0 and (cfg["coarse"] or cfg["fine"])) else torch.zeros)(
            N, gemb, dtype=torch.float32, device=dev)
        wc = _weights_struct(pc, depth, W, temb, gemb, cfg["enabled"], cfg["precision"]) if cfg["coarse"] else None
        wf = _weights_struct(pf, depth, W, temb, gemb, cfg["enabled"], cfg["precision"]) if cfg["fine"] else None
        sgc = _grads_struct(gc, depth, cfg["enabled"]) if cfg["coarse"] else None
        sgf = _grads_struct(gf, depth, cfg["enabled"]) if cfg["fine"] else None
        ptr = lambda x: x.data_ptr() if x is not None else None  # noqa: E731
        side = _overlap["stream"] if N > 0 else None
        k_out, k_sub = d_out, d_sub
        if side is not None:
            # the caller all-reduces the pass-through gradients in place while these kernels run: they read copies
            if _overlap["clone"]:
                k_out = [None if x is None else x.clone() for x in d_out]
                k_sub = [None if x is None else x.clone() for x in d_sub]
            side.wait_stream(torch.cuda.current_stream(dev))
            if _overlap["sm_reserve"]:
                with torch.cuda.device(dev):     # the reserve is kept per device (the one current at the call)
                    lib.ed3dgs_deform_set_sm_reserve(_overlap["sm_reserve"])
        with torch.cuda.device(dev), torch.cuda.stream(side):
            if N > 0:
                d_out_keep, d_sub_keep = d_out, d_sub
                d_out, d_sub = k_out, k_sub
                dout_arr = abi.PtrArr5(*[ptr(x) for x in d_out])
                dsub_arr = abi.PtrArr5(*[ptr(x) for x in d_sub])
                coef = abi.FloatArr5(*cfg["coef"])
                _lib.check(lib.ed3dgs_deform_backward(
                    ctypes.byref(wc) if wc else None, ctypes.byref(wf) if wf else None,
                    ctypes.byref(sgc) if sgc else None, ctypes.byref(sgf) if sgf else None, N, emb_c.data_ptr(),
                    tau_c.data_ptr(), tau_f.data_ptr(), ctypes.byref(dout_arr), ctypes.byref(dsub_arr),
                    ctypes.byref(coef), work.data_ptr(), work.numel(), d_emb.data_ptr(), d_tau_c.data_ptr(),
                    d_tau_f.data_ptr(), _lib.stream_handle()))
                max_emb, dim = table_c.shape
                for on, rows, d_tau in ((cfg["coarse"], cfg["rows_c"], d_tau_c), (cfg["fine"], cfg["rows_f"], d_tau_f)):
                    if on:
                        _lib.check(lib.ed3dgs_temporal_embed_backward(
                            table_c.data_ptr(), max_emb, dim, int(rows), t_c.data_ptr(), d_tau.data_ptr(),
                            d_table.data_ptr(), d_t.data_ptr(), _lib.stream_handle()))
                d_out, d_sub = d_out_keep, d_sub_keep
                if ctx.needs_input_grad[2]:   # the camera's time offset (same stream as the kernels)
                    cam_no = cfg["cam_no"]
                    if cam_no is not None:
                        d_offsets[cam_no] = d_t
                    else:   # offset = mean of the non-zero rows; no non-zero row: the reference's nan -> 0, no gradient
                        nz = offsets.ne(0)
                        d_offsets.copy_(torch.where(nz, d_t / nz.sum().clamp_min(1), torch.zeros_like(d_t)).view(offsets.shape))
        if side is not None:
            if _overlap["sm_reserve"]:
                with torch.cuda.device(dev):
                    lib.ed3dgs_deform_set_sm_reserve(0)
            # (tested before this backward is parked: a non-empty `keep` means an earlier backward of the same step)
            serial = _must_serialise(getattr(ctx, "leaf_refs", ()))
            # (not flat / d_emb: a second reference would make AccumulateGrad copy them -- on the main stream, too early)
            _overlap["keep"].append((k_out, k_sub, work, t_c, table_c, emb_c, tau_c, tau_f, offsets, params))
            if serial:
                _overlap["serialised"] += 1
            if serial or ctx.needs_input_grad[1]:   # (the time's own gradient is consumed on the main stream too)
                torch.cuda.current_stream(dev).wait_stream(side)

        def passthrough(h):
            a, b = d_out[h], d_sub[h]
            if a is None and b is None:
                return None
            g = a if b is None else (b if a is None else a + b)
            return g.view(N, 16, 3) if h == 4 else g

        d_ins = [passthrough(h) for h in range(5)]
        return (None, d_t.view(()).reshape(ctx.cfg["t_shape"]) if ctx.needs_input_grad[1] else None,
                d_offsets if ctx.needs_input_grad[2] else None, d_table, d_emb, *d_ins, *gc, *gf)


class deform_network(nn.Module):
    def __init__(self, D=8, W=256, min_embeddings=30, max_embeddings=150, num_frames=300, num_cam=None, args=None):
        super().__init__()
        self.D = D
        self.W = W
        self.args = args
        self.min_embeddings = min_embeddings
        self.max_embeddings = max_embeddings
        self.num_frames = num_frames
        self.temporal_embedding_dim = args.temporal_embedding_dim
        self.gaussian_embedding_dim = args.gaussian_embedding_dim
        self.c2f_temporal_iter = args.c2f_temporal_iter
        # "fp32" (default): 3xTF32 tensor-core GEMMs, fp32-grade parity with the reference's cuBLAS SGEMM path;
        # "tf32": one tf32 MMA per product (torch's allow_tf32 arithmetic), documented looser tolerance
        self.precision = getattr(args, "mlp_precision", "fp32")

        (self.feature_out_c, self.pos_deform_c, self.scales_deform_c, self.rotations_deform_c,
         self.opacity_deform_c, self.rgb_deform_c) = self.create_net()
        (self.feature_out_f, self.pos_deform_f, self.scales_deform_f, self.rotations_deform_f,
         self.opacity_deform_f, self.rgb_deform_f) = self.create_net()

        if args.zero_temporal:
            self.weight = torch.nn.Parameter(torch.zeros(max_embeddings, self.temporal_embedding_dim))
        else:
            self.weight = torch.nn.Parameter(torch.normal(
                0., 0.01 / np.sqrt(self.temporal_embedding_dim), size=(max_embeddings, self.temporal_embedding_dim)))
        self.offsets = torch.nn.Parameter(torch.zeros((30, 1)))

    def create_net(self):
        layers = [nn.Linear(self.temporal_embedding_dim + self.gaussian_embedding_dim, self.W)]
        for _ in range(self.D - 1):
            layers.append(nn.ReLU())
            layers.append(nn.Linear(self.W, self.W))
        head = lambda k: nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, k))  # noqa: E731
        return nn.Sequential(*layers), head(3), head(3), head(4), head(1), head(3 * 16)

    def int_lininterp(self, t, init_val, final_val, until):
        return int(init_val + (final_val - init_val) * min(max(t, 0), until) / until)

    def get_temporal_embed(self, t, current_num_embeddings, align_corners=True):
        """[N, dim] like the reference (the row is computed once and expanded)."""
        N = t.shape[0]
        tau = temporal_embed(self.weight.detach().contiguous(), current_num_embeddings,
                             t[0, 0].detach().reshape(1).float().contiguous())
        return tau.unsqueeze(0).expand(N, -1)

    def _rows(self, iter, num_down_emb, use_coarse):
        if use_coarse:
            return num_down_emb
        if self.args.no_c2f_temporal_embedding:
            return self.max_embeddings
        return self.int_lininterp(iter, num_down_emb, self.max_embeddings, self.c2f_temporal_iter)

    def forward(self, point, scales=None, rotations=None, opacity=None, time_emb=None, cam_no=None, pc=None,
                embeddings=None, sh_coefs=None, iter=None, num_down_emb_c=30, num_down_emb_f=30):
        pts, scales, rotations, opacity = point[:, :3], scales[:, :3], rotations[:, :4], opacity[:, :1]
        orig = (pts, scales, rotations, opacity, sh_coefs)
        a = self.args
        with torch.no_grad():   # the gradient of the offsets comes out of the autograd node below
            if cam_no is None:
                offset = torch.masked_select(self.offsets, self.offsets.ne(0)).mean()
                offset[torch.isnan(offset)] = 0
            else:
                offset = self.offsets[cam_no]
            time_emb += offset

        use_anneal = a.use_anneal
        coef = 1 if not use_anneal else np.clip(iter / 1000, 0, 1)
        coef_c = 1 if not use_anneal else np.clip((iter - a.deform_from_iter) / 1000, 0, 1)
        coef_o = coef_c
        coef_s = coef_c
        coarse_on, fine_on = not a.no_coarse_deform, not a.no_fine_deform
        if not coarse_on and not fine_on:
            return (*orig, (orig, orig))

        emb = pc.get_embedding if pc is not None else embeddings
        enabled = (True, not a.no_ds, not a.no_dr, not a.no_do, not a.no_dc)
        if sh_coefs is None:
            if enabled[4]:
                raise _lib.Ed3dgsError("sh_coefs is required unless no_dc is set")
            sh_in = torch.zeros(pts.shape[0], 16, 3, device=pts.device)
        else:
            sh_in = sh_coefs
        heads_c = (self.pos_deform_c, self.scales_deform_c, self.rotations_deform_c, self.opacity_deform_c,
                   self.rgb_deform_c)
        heads_f = (self.pos_deform_f, self.scales_deform_f, self.rotations_deform_f, self.opacity_deform_f,
                   self.rgb_deform_f)
        params_c, depth = _stage_tensors(self.feature_out_c, heads_c)
        params_f, _ = _stage_tensors(self.feature_out_f, heads_f)
        t = time_emb[0, :1]
        cfg = dict(W=self.W, depth=depth, temb=self.temporal_embedding_dim, gemb=self.gaussian_embedding_dim,
                   enabled=enabled, coarse=coarse_on, fine=fine_on, n_params_c=len(params_c),
                   rows_c=self._rows(iter, num_down_emb_c, a.use_coarse_temporal_embedding),
                   rows_f=self._rows(iter, num_down_emb_f, False),
                   coef=(float(coef), float(coef * coef_s), float(coef), float(coef * coef_o), float(coef_c)),
                   t_shape=tuple(t.shape), precision=_PRECISIONS[self.precision], cam_no=cam_no)
        res = _DeformFunction.apply(cfg, t, self.offsets, self.weight, emb, pts, scales, rotations, opacity, sh_in,
                                    *params_c, *params_f)
        sub, out = list(res[:5]), list(res[5:])
        if sh_coefs is None:
            sub[4], out[4] = None, None
        return (*out, (tuple(sub), orig))

    def get_mlp_parameters(self):
        return [param for name, param in self.named_parameters() if name != "offsets"]


def initialize_weights(m):
    if isinstance(m, nn.Linear):
        nn.init.xavier_uniform_(m.weight, gain=1)
        if m.bias is not None:
            nn.init.xavier_uniform_(m.weight, gain=1)
