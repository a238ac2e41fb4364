"""The per-iteration update of RAFT and the whole test-mode forward around the correlation path (SURVEY.md section 8,
rows f1-f3): host-side mirrors of the reference's `BasicUpdateBlock` / `SmallUpdateBlock` (core/update.py:99-136) and of
`RAFT.forward(..., test_mode=True)` (core/raft.py:86-144) that run a model the USER built from the reference's own
classes -- its weights, its encoders -- with the hot layers on this library's kernels:

  lookup -> convc1 -> ReLU          CorrBlock.lookup_convc1            (rc_corr_lookup_convc1_fwd)
  convc2 / convf2 / conv            conv2d_fused                       (rc_conv2d_ex_fwd; `conv` writes straight into the
                                                                        GRU's input tensor, no torch.cat)
  GRU half (z, r, r*h, q, blend)    two conv2d_fused launches          (epilogues "mul_from" and "gru")
  flow head, mask head's 3x3        conv2d_fused
  encoder output layer              encoder_head                       (rc_fnet_conv1x1_prep)
  convex upsampling                 upsample_flow                      (rc_convex_upsample_fwd)

  convf1 (7x7 on the flow) + ReLU    small_conv_relu                    (rc_small_conv_relu_fwd)
  coords1 += delta; flow = ...       flow_update                        (rc_flow_update)

What stays torch / cuDNN: the encoders' residual stages and the mask head's 576-wide 1x1.
In test mode only the last iteration's upsampled flow is returned (core/raft.py:141-142), so the mask head and the
upsampling run once, after the loop.  Nothing here imports the reference: it duck-types the modules it is given.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .corr import AlternateCorrBlock, CorrBlock, encoder_head
from .utils import conv2d_fused, coords_grid, flow_update, small_conv_relu, upsample_flow

__all__ = ["UpdateBlock", "FusedRAFT"]


class _Stacked:
    """torch.cat of parameter tensors, redone when one of them changes (identity or in-place version)."""

    def __init__(self):
        self._d = {}

    def get(self, name, tensors):
        key = tuple((id(t), t._version) for t in tensors)
        ent = self._d.get(name)
        if ent is None or ent[0] != key:
            ent = self._d[name] = (key, torch.cat([t.detach() for t in tensors]).contiguous(), tensors)
        return ent[1]


class UpdateBlock:
    """`BasicUpdateBlock` / `SmallUpdateBlock` of the reference (core/update.py:99-136), inference only.

    ub: a module with the reference's attributes (`encoder.convc1 ... conv`, `gru.convz1 ...` or `gru.convz ...`,
    `flow_head.conv1/conv2`, optionally `mask`).  `__call__(net, inp, corr, flow)` has the reference's signature and
    results `(net, mask, delta_flow)`; `step(...)` is the same iteration with the lookup fused into convc1."""

    def __init__(self, ub):
        self.ub = ub
        self._stk = _Stacked()
        self.has_mask = hasattr(ub, "mask")
        self.sep = hasattr(ub.gru, "convz1")

    # ---- motion encoder (core/update.py:71-77, 90-97) -> the GRU input x = [inp | out | flow]; x's flow channels
    # already hold `flow` (dense copy: the input of convf1)
    def _motion_into_x(self, x, cdim, cor, flow):
        me = self.ub.encoder
        if hasattr(me, "convc2"):
            cor = conv2d_fused(cor, me.convc2.weight, me.convc2.bias, "relu")
        k = me.convf1.weight.shape[-1]
        flo = small_conv_relu(flow, me.convf1.weight, me.convf1.bias) if k <= 7 and flow.shape[1] <= 4 else F.relu(me.convf1(flow))
        flo = conv2d_fused(flo, me.convf2.weight, me.convf2.bias, "relu")
        co = me.conv.weight.shape[0]
        conv2d_fused((cor, flo), me.conv.weight, me.conv.bias, "relu", out=x[:, cdim:cdim + co])  # no torch.cat: written in place
        return x

    def flow_slice(self, x, cdim):
        co = self.ub.encoder.conv.weight.shape[0]
        return x[:, cdim + co:cdim + co + 2]

    def _gru_half(self, h, x, cz, cr, cq, name):
        hd = h.shape[1]
        wzr = self._stk.get(name + "w", (cz.weight, cr.weight))
        bzr = self._stk.get(name + "b", (cz.bias, cr.bias))
        zr = conv2d_fused((h, x), wzr, bzr, "sigmoid", epilogue="mul_from", split=hd, aux1=h)  # [z | r*h]
        return conv2d_fused((zr[:, hd:], x), cq.weight, cq.bias, "tanh", epilogue="gru", aux1=zr[:, :hd], aux2=h)

    def _gru(self, h, x):
        g = self.ub.gru
        if self.sep:  # SepConvGRU: horizontal then vertical (core/update.py:33-60)
            h = self._gru_half(h, x, g.convz1, g.convr1, g.convq1, "1")
            return self._gru_half(h, x, g.convz2, g.convr2, g.convq2, "2")
        return self._gru_half(h, x, g.convz, g.convr, g.convq, "0")  # ConvGRU (:16-31)

    def _x_buffer(self, inp, flow):
        me = self.ub.encoder
        B, cdim, H, W = inp.shape
        x = torch.empty((B, cdim + me.conv.weight.shape[0] + 2, H, W), dtype=torch.float32, device=inp.device)
        x[:, :cdim] = inp
        return x

    def flow_delta(self, net):
        fh = self.ub.flow_head  # core/update.py:6-14
        return conv2d_fused(conv2d_fused(net, fh.conv1.weight, fh.conv1.bias, "relu"), fh.conv2.weight, fh.conv2.bias)

    def up_mask(self, net):
        if not self.has_mask:
            return None
        m = conv2d_fused(net, self.ub.mask[0].weight, self.ub.mask[0].bias, "relu")  # core/update.py:122-125
        return 0.25 * self.ub.mask[2](m)                                               # :134-135

    def heads(self, net, want_mask=True):
        return (self.up_mask(net) if want_mask else None), self.flow_delta(net)

    def __call__(self, net, inp, corr, flow, upsample=True):
        """The reference's signature: corr is the lookup result (B, L*(2r+1)^2, H, W)."""
        with torch.no_grad():
            me = self.ub.encoder
            cor = conv2d_fused(corr.float().contiguous(), me.convc1.weight, me.convc1.bias, "relu")
            flow = flow.float().contiguous()
            x = self._x_buffer(inp.float(), flow)
            self.flow_slice(x, inp.shape[1]).copy_(flow)
            x = self._motion_into_x(x, inp.shape[1], cor, flow)
            net = self._gru(net.float().contiguous(), x)
            mask, delta = self.heads(net, upsample)
            return net, mask, delta

    def step(self, net, x, cdim, corr_fn, coords1, flow, want_mask=False):
        """One iteration with the lookup fused into convc1; x is the persistent GRU input whose first cdim channels
        hold `inp` (see `FusedRAFT`)."""
        me = self.ub.encoder
        if isinstance(corr_fn, CorrBlock) and corr_fn._tiled is not None:
            cor = corr_fn.lookup_convc1(coords1, me.convc1.weight, me.convc1.bias)
        else:  # on-the-fly block (or a row-layout one): plain lookup, convc1 as a 1x1 convolution
            cor = conv2d_fused(corr_fn(coords1), me.convc1.weight, me.convc1.bias, "relu")
        self._motion_into_x(x, cdim, cor, flow)
        net = self._gru(net, x)
        mask, delta = self.heads(net, want_mask)
        return net, mask, delta


class FusedRAFT:
    """`RAFT.forward(image1, image2, iters, flow_init, upsample, test_mode=True)` (core/raft.py:86-144) for a RAFT
    module built from the reference's classes, with the correlation path, the update block and the upsampling on this
    library's kernels.  graph=True captures the whole forward in a CUDA graph per (image shape, iters, flow_init given)
    and replays it: what the ROS node's callback needs (ros/scripts/main.py:117-149, one pair per message).

    Returns (flow_low, flow_up) like the reference's test mode.  Training: use the reference's forward with
    `core.raft.CorrBlock = raft_ros_b200.CorrBlock`."""

    def __init__(self, model, precision="bf16x3", graph=False):
        self.model = model
        self.precision = precision
        self.graph = graph
        self.update = UpdateBlock(model.update_block)
        self._graphs = {}
        self._side = {}

    def _side_stream(self, device):
        s = self._side.get(device.index)
        if s is None:
            s = self._side[device.index] = torch.cuda.Stream(device)
        return s

    def _forward(self, image1, image2, iters, flow_init):
        m = self.model
        args = m.args
        image1 = (2 * (image1 / 255.0) - 1.0).contiguous()
        image2 = (2 * (image2 / 255.0) - 1.0).contiguous()
        hdim, cdim = m.hidden_dim, m.context_dim
        amp = bool(getattr(args, "mixed_precision", False))
        enc = m.fnet
        # the context encoder does not depend on the feature encoder (core/raft.py:99-114): it runs on a second stream
        cur = torch.cuda.current_stream(image1.device)
        side = self._side_stream(image1.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.autocast("cuda", enabled=amp):
            cnet = m.cnet(image1)
        with torch.autocast("cuda", enabled=amp):
            y = torch.cat([image1, image2], dim=0)  # core/extractor.py:168-192 (small :239-264) up to conv2's input
            y = enc.layer3(enc.layer2(enc.layer1(enc.relu1(enc.norm1(enc.conv1(y))))))
        if getattr(args, "alternate_corr", False):
            f = enc.conv2(y.float()).float()
            corr_fn = AlternateCorrBlock(f[:f.shape[0] // 2], f[f.shape[0] // 2:], radius=args.corr_radius, precision=self.precision)
        else:
            prep = encoder_head(y.float(), enc.conv2.weight, enc.conv2.bias)  # conv2 + .float() + the build's layout prepass
            corr_fn = CorrBlock.from_prepared(prep, radius=args.corr_radius, precision=self.precision)
        cur.wait_stream(side)
        cnet.record_stream(cur)
        net, inp = torch.split(cnet.float().contiguous(), [hdim, cdim], dim=1)
        net, inp = torch.tanh(net).contiguous(), torch.relu(inp)
        N, _, H, W = image1.shape
        coords0 = coords_grid(N, H // 8, W // 8, device=image1.device)
        coords1 = coords0.clone()
        if flow_init is not None:
            coords1 = coords1 + flow_init
        flow = coords1 - coords0
        x = self.update._x_buffer(inp, flow)
        xflow = self.update.flow_slice(x, cdim)
        xflow.copy_(flow)
        for it in range(iters):
            net, _, delta = self.update.step(net, x, cdim, corr_fn, coords1, flow)
            flow_update(coords1, delta, coords0, flow, xflow)  # coords1 += delta; flow = coords1 - coords0 (core/raft.py:123,131-132)
        if self.update.has_mask:
            # the mask of the last iteration's hidden state (core/raft.py:128,137)
            flow_up = upsample_flow(flow, self.update.up_mask(net))
        else:  # raft-small: upflow8 (core/utils/utils.py:80-82)
            flow_up = 8 * F.interpolate(flow, size=(8 * flow.shape[2], 8 * flow.shape[3]), mode="bilinear", align_corners=True)
        return flow, flow_up

    def __call__(self, image1, image2, iters=12, flow_init=None, upsample=True, test_mode=True):
        if not test_mode:
            raise RuntimeError("FusedRAFT is the test-mode forward; train through the reference's forward with the drop-in CorrBlock")
        with torch.no_grad():
            if not self.graph:
                return self._forward(image1, image2, iters, flow_init)
            # a graph holds the converted weights it was captured with: weights changed in place since (load_state_dict,
            # an optimiser step) invalidate every captured graph
            stamp = sum(p._version for p in self.model.parameters())
            if stamp != getattr(self, "_stamp", None):
                self._graphs.clear()
                self._stamp = stamp
            key = (tuple(image1.shape), image1.device.index, int(iters), flow_init is not None)
            ent = self._graphs.get(key)
            if ent is None:
                s1, s2 = image1.detach().float().clone(), image2.detach().float().clone()
                sf = None if flow_init is None else flow_init.detach().float().clone()
                for _ in range(2):  # warm-up outside the capture: weight conversions, lazy module loading, cuDNN plans
                    self._forward(s1, s2, iters, sf)
                torch.cuda.synchronize(image1.device)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    out = self._forward(s1, s2, iters, sf)
                ent = self._graphs[key] = (g, s1, s2, sf, out)
            g, s1, s2, sf, out = ent
            s1.copy_(image1)
            s2.copy_(image2)
            if sf is not None:
                sf.copy_(flow_init)
            g.replay()
            return out[0].clone(), out[1].clone()
