"""torch.autograd bridges onto the C ABI (include/b200lmsfc.h).

PyTorch supplies device memory, streams and the autograd tape; all arithmetic happens in
libb200lmsfc.so.  Layout contract: activations are bf16 tensors of logical shape [B,C,H,W] stored
channels-last (NHWC in memory); latents and likelihoods are fp32 channels-last.

Parameter gradients are written by the backward kernels straight into `param.grad` (pre-allocated
by the model as views of one flat arena, see model.ParamArena); the autograd Functions therefore
return None for parameters.  `grad_ready` callbacks let the data-parallel wrapper launch bucketed
all-reduces while backward is still running.
"""
import math

import torch

from . import _lib as L
from ._lib import ConvDesc, ptr, stream

BF16, F32 = torch.bfloat16, torch.float32
CL = torch.channels_last


class SideStreams:
    """Concurrency inside one train step (CUDA streams forked / joined inside the captured graph):

    * `wgrad`: weight-gradient GEMMs (+ bias column sums) are off the backward critical path -- only the
      optimiser needs them -- so they run beside the data-gradient chain; the latency-bound latent-resolution
      layers (a few CTAs each) overlap almost entirely.
    * `hyper`: in noise mode y_hat = y + U does not depend on the hyperprior, so h_a -> EntropyBottleneck ->
      h_s -> Gaussian likelihood runs beside the synthesis transform (and its backward beside g_s's backward;
      autograd replays every node on the stream it was recorded on).
    * `opt`: the pipelined train step (train.TrainStep.capture(pipelined=True)) runs clip + Adam of the PREVIOUS
      backward pass here, slice by slice in the order the forward pass consumes the weights, so the HBM-bound
      update hides under the tensor-bound forward convolutions.
    * `heads`: the 3x3 output heads of g_s (p5, p4, p3 reconstructions) only feed the distortion term, so they run
      beside the next up-sampling stage; in backward their data gradients run beside the up-sampling stage's.
    Disabled (None) by default: stand-alone modules run on the caller's stream."""
    N_WGRAD = int(__import__("os").environ.get("LMSFC_N_WGRAD", "3"))
    _wgrad_pool = []
    _wgrad_next = 0
    wgrad = None
    hyper = None
    opt = None
    heads = None
    _join_queued = False

    @classmethod
    def enable(cls, device, wgrad=True, hyper=True, opt=True, heads=True):
        cls.wgrad = torch.cuda.Stream(device=device) if wgrad else None
        n = int(wgrad) if wgrad is not True else cls.N_WGRAD
        cls._wgrad_pool = [cls.wgrad] + [torch.cuda.Stream(device=device) for _ in range(max(0, n - 1))] if wgrad else []
        cls._wgrad_next = 0
        cls.hyper = torch.cuda.Stream(device=device, priority=-1) if hyper else None   # on the forward's tail: A/B 1.036 -> 1.014 ms
        cls.opt = torch.cuda.Stream(device=device) if opt else None
        cls.heads = torch.cuda.Stream(device=device, priority=-1) if heads else None   # their outputs gate the loss: A/B 1.008 -> 1.000 ms

    @classmethod
    def disable(cls):
        cls.wgrad = cls.hyper = cls.opt = cls.heads = None
        cls._wgrad_pool = []

    @classmethod
    def next_wgrad(cls):
        """Round-robin over the weight-gradient streams: consecutive layers' (latency-bound) gradient GEMMs overlap."""
        if not cls._wgrad_pool:
            return None
        s = cls._wgrad_pool[cls._wgrad_next % len(cls._wgrad_pool)]
        cls._wgrad_next += 1
        return s

    @classmethod
    def join_wgrad(cls):
        cur = torch.cuda.current_stream()
        for s in cls._wgrad_pool:
            cur.wait_stream(s)

    @classmethod
    def queue_join(cls):
        """Join the weight-gradient stream at the end of the running backward pass."""
        if not cls._join_queued:
            cls._join_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(cls._join)

    @classmethod
    def _join(cls):
        cls._join_queued = False
        cls.join_wgrad()


def _require_cuda(t):
    if not t.is_cuda:
        raise RuntimeError("l-msfc-v2-training_b200 runs on CUDA tensors only (no CPU fallback)")


def nhwc(t, dtype):
    """Return `t` ([B,C,H,W]) as a channels-last tensor of `dtype` (no copy when already so)."""
    _require_cuda(t)
    if t.dim() != 4:
        raise ValueError(f"expected a 4-D [B,C,H,W] tensor, got {tuple(t.shape)}")
    if t.dtype == dtype and t.is_contiguous(memory_format=CL):
        return t
    if t.dtype == F32 and dtype == BF16 and t.is_contiguous() and not t.requires_grad:
        B, C, H, W = t.shape
        out = torch.empty((B, C, H, W), dtype=BF16, device=t.device, memory_format=CL)
        L.nchw_to_nhwc_bf16(ptr(t), ptr(out), B, C, H * W, stream())
        return out
    return t.to(dtype=dtype, memory_format=CL)


def empty_nhwc(B, C, H, W, dtype, device):
    return torch.empty((B, C, H, W), dtype=dtype, device=device, memory_format=CL)


def grad_view(param, layout_perm=None):
    """`param.grad` (allocated zero-filled with the parameter's own strides when missing) and,
    for permuted GEMM-layout parameters, its physically contiguous view."""
    if param.grad is None:
        param.grad = torch.zeros_like(param)   # preserve_format keeps the GEMM-layout strides
    g = param.grad
    if layout_perm is not None:
        g = g.permute(*layout_perm)
    if not g.is_contiguous():
        return None
    return g


# ====================================================================== conv / deconv
class ConvState:
    """Per-layer device state shared by forward and backward: geometry and the bf16 weight shadow
    ([D0][T][D1]; read K-major by conv-forward / deconv-dgrad and MN-major by conv-dgrad /
    deconv-forward, so there is no transposed copy)."""

    def __init__(self, kind, k, stride, c0, c1, cout):
        self.kind, self.k, self.stride, self.c0, self.c1, self.cout = kind, k, stride, c0, c1, cout
        self.T = k * k
        cin = c0 + c1
        self.D0, self.D1 = (cout, cin) if kind == 0 else (cin, cout)
        self.w_direct = None
        self._version = -1
        self.grad_ready = None     # callable(list_of_params) set by the DP wrapper

    def out_hw(self, H, W, target=None):
        if self.kind == 0:
            return (H + self.stride - 1) // self.stride, (W + self.stride - 1) // self.stride
        if target is None:
            return H * self.stride, W * self.stride
        return int(target[0]), int(target[1])

    def desc(self, B, Hin, Win, Hout, Wout):
        return ConvDesc(self.kind, self.k, self.stride, B, Hin, Win, Hout, Wout, self.c0, self.c1, self.cout)

    def gemm_weight(self, weight):
        """[D0][T][D1]-contiguous view of the torch-layout parameter ([D0][D1][kh][kw])."""
        w = weight.detach().permute(0, 2, 3, 1)
        return w if w.is_contiguous() else None

    def refresh(self, weight, force=False):
        """(Re)build the bf16 shadow when the fp32 parameter changed."""
        v = weight._version
        if not force and self.w_direct is not None and v == self._version:
            return
        wg = self.gemm_weight(weight)
        if wg is None:   # parameter not in GEMM layout (user-assigned tensor): stage a permuted copy
            wg = weight.detach().permute(0, 2, 3, 1).contiguous()
        if self.w_direct is None:
            self.w_direct = torch.empty((self.D0, self.T, self.D1), dtype=BF16, device=weight.device)
        st = stream()
        L.cast_f32_bf16(ptr(wg), ptr(self.w_direct), wg.numel(), st)
        self._version = v


class GDNLink:
    """What a GDN layer shares with the consumers of its output y during one forward / backward pass, so that the
    consumer's data-gradient kernel can run the GDN's backward contraction in its own epilogue
    (lmsfc_conv_dgrad_gdn_backward): the GDN's saved x, rn, gamma, and -- in backward -- the gradient of a second
    ("side") consumer of y, which the main consumer's kernel adds before the contraction.

    Protocol (all decided while the forward pass is recorded, so a captured graph replays it):
      * the MAIN consumer's forward sets `main_fuses` (its dgrad launch is fusable: csrc/igemm.cu cost model);
      * a SIDE consumer that runs after it and sees main_fuses defers: in backward it parks its data gradient in `addend`
        (returning None to autograd) instead of letting autograd add the two gradients with a separate kernel;
        a side consumer that ran BEFORE the main one registers `side_plain`, and the main consumer then does not fuse;
      * autograd runs the later-recorded node first, so the model records main before side (model.LMSFC.synthesis).
    """
    __slots__ = ("x", "rn", "gamma_bf", "inverse", "main_fuses", "side_plain", "n_side", "addend", "addend_ev", "t",
                 "fused_done", "expect_side")

    def __init__(self, x, rn, gamma_bf, inverse):
        self.x, self.rn, self.gamma_bf, self.inverse = x, rn, gamma_bf, bool(inverse)
        self.main_fuses = None
        self.expect_side = False       # set by the model when y has a second consumer (recorded after the main one)
        self.side_plain = False
        self.n_side = 0
        self.addend = self.addend_ev = self.t = None
        self.fused_done = False

    def register_consumer(self, state, d, side):
        """Called from the consumer's forward; returns the consumer's role: "main", "side" or None (plain)."""
        if side:
            if self.main_fuses:
                self.n_side += 1
                if self.n_side > 1:
                    raise RuntimeError("GDNLink: one deferred side consumer per GDN output")
                return "side"
            self.side_plain = True
            return None
        if self.main_fuses is not None:
            raise RuntimeError("GDNLink: a GDN output has one main consumer")
        self.main_fuses = (not self.side_plain and (not self.expect_side or (L.FUSE_MODE & 4) != 0)
                           and state.c0 == self.x.shape[1] == 256 and L.gdn_fusable(d, dgrad=True, n_in=state.c0))
        return "main" if self.main_fuses else None


def _conv_backward(state, shape, x0, x1, g, weight, bias, need_dx, link, role):
    """Data gradient (critical path) then weight / bias gradients (side stream) of one conv / deconv.
    With a GDNLink the data gradient either carries the producing GDN's backward (role "main": returns the gradient
    w.r.t. the GDN's INPUT) or is parked for the main consumer (role "side": returns None)."""
    B, H, W, Ho, Wo = shape
    st = stream()
    d = state.desc(B, H, W, Ho, Wo)
    gx = None
    if need_dx:
        gx = empty_nhwc(B, state.c0, H, W, BF16, g.device)
        if role == "main":
            if link.n_side and link.addend is None:
                raise RuntimeError("GDNLink: the side consumer's backward has not run yet (autograd order changed?)")
            if link.addend_ev is not None:
                torch.cuda.current_stream().wait_event(link.addend_ev)
                link.addend.record_stream(torch.cuda.current_stream())
            t = torch.empty_like(link.x)
            L.conv_dgrad_gdn_backward(d, ptr(g), ptr(state.w_direct), state.c0, ptr(link.addend), ptr(link.x), ptr(link.rn),
                                      ptr(link.gamma_bf), int(link.inverse), ptr(t), ptr(gx), st)
            link.t, link.fused_done = t, True
        else:
            L.conv_dgrad(d, ptr(g), ptr(state.w_direct), state.c0, ptr(gx), st)
            if role == "side":
                ev = torch.cuda.Event()
                ev.record()
                link.addend, link.addend_ev = gx, ev
                gx = None
    if weight.requires_grad:
        ws = SideStreams.next_wgrad()
        if ws is not None:
            ev = torch.cuda.Event()
            ev.record()                      # g is complete on the current stream
            for t_ in (g, x0, x1):
                if t_ is not None:
                    t_.record_stream(ws)
            with torch.cuda.stream(ws):
                ws.wait_event(ev)
                _conv_wgrad(state, d, x0, x1, g, weight, bias)
            SideStreams.queue_join()
        else:
            _conv_wgrad(state, d, x0, x1, g, weight, bias)
    return gx


def _conv_wgrad(state, d, x0, x1, g, weight, bias):
    st = stream()
    ready = []
    gw = grad_view(weight, (0, 2, 3, 1))
    gb = grad_view(bias) if (bias is not None and bias.requires_grad) else None
    if gw is not None:
        L.conv_wgrad(d, ptr(x0), ptr(x1), ptr(g), ptr(gw), ptr(gb), st)
    else:  # foreign (non GEMM-layout) parameter: stage and add
        tmp = torch.zeros((state.D0, state.T, state.D1), dtype=F32, device=g.device)
        L.conv_wgrad(d, ptr(x0), ptr(x1), ptr(g), ptr(tmp), ptr(gb), st)
        weight.grad.add_(tmp.view(state.D0, state.k, state.k, state.D1).permute(0, 3, 1, 2))
    ready.append(weight)
    if gb is not None:
        ready.append(bias)
    if state.grad_ready is not None:
        state.grad_ready(ready)        # the DP reducer records its event on the stream the gradient was produced on


class _ConvFn(torch.autograd.Function):
    """conv / deconv (+bias, +ReLU / LeakyReLU) -- compressai/models/utils.py:128-146.
    relu: 0 none, 1 ReLU, 2 LeakyReLU(0.01).  mode 0: returns bf16 activation; mode 1 (latent head): returns (fp32 y,
    bf16 |y|); mode 2 (latent head of the mean-scale hyperprior, models/google.py:395-398): returns (fp32 y, bf16 y).
    `link` (GDNLink of the GDN that produced x0) / `side`: see GDNLink."""

    @staticmethod
    def forward(ctx, x0, x1, weight, bias, state, relu, mode, out_hw, link, side):
        B, _, H, W = x0.shape
        Ho, Wo = state.out_hw(H, W, out_hw)
        state.refresh(weight)
        d = state.desc(B, H, W, Ho, Wo)
        w = state.w_direct
        dev = x0.device
        out_b = out_f = out_a = None
        if mode == 0:
            out_b = empty_nhwc(B, state.cout, Ho, Wo, BF16, dev)
        elif mode == 1:
            out_f = empty_nhwc(B, state.cout, Ho, Wo, F32, dev)
            out_a = empty_nhwc(B, state.cout, Ho, Wo, BF16, dev)
        else:
            out_f = empty_nhwc(B, state.cout, Ho, Wo, F32, dev)
            out_b = empty_nhwc(B, state.cout, Ho, Wo, BF16, dev)
        L.conv_forward(d, ptr(x0), ptr(x1), ptr(w), ptr(bias), int(relu), ptr(out_b), ptr(out_f), ptr(out_a), stream())
        ctx.state, ctx.relu, ctx.mode, ctx.shape = state, relu, mode, (B, H, W, Ho, Wo)
        ctx.weight, ctx.bias = weight, bias
        ctx.need_dx = ctx.needs_input_grad[0]
        ctx.link = link
        ctx.role = link.register_consumer(state, d, side) if (link is not None and ctx.need_dx) else None
        if mode == 0:
            ctx.save_for_backward(x0, x1, out_b if relu else None, None)
            return out_b
        ctx.save_for_backward(x0, x1, None, out_f)
        return out_f, (out_a if mode == 1 else out_b)

    @staticmethod
    def backward(ctx, *grads):
        state = ctx.state
        x0, x1, out_relu, y_f32 = ctx.saved_tensors
        B, H, W, Ho, Wo = ctx.shape
        st = stream()
        if ctx.mode == 0:
            g = nhwc(grads[0], BF16)
            if ctx.relu:
                g = g.clone(memory_format=CL)
                if ctx.relu == 2:
                    L.leaky_relu_backward(ptr(g), ptr(out_relu), g.numel(), LEAKY_SLOPE, st)
                else:
                    L.relu_backward(ptr(g), ptr(out_relu), g.numel(), st)
        else:
            g_direct, g_abs = grads
            g = empty_nhwc(B, state.cout, Ho, Wo, BF16, x0.device)
            gd = None if g_direct is None else nhwc(g_direct, F32)
            ga = None if g_abs is None else nhwc(g_abs, BF16)
            if ctx.relu == 2:
                raise NotImplementedError("a latent head ends in no activation or ReLU (models/google.py:275-282)")
            L.abs_backward(ptr(gd), ptr(ga), ptr(y_f32), g.numel(), int(ctx.relu) | (2 if ctx.mode == 2 else 0), ptr(g), st)
        # only source 0 is differentiable: source 1 is input data
        gx = _conv_backward(state, ctx.shape, x0, x1, g, ctx.weight, ctx.bias, ctx.need_dx, ctx.link, ctx.role)
        return gx, None, None, None, None, None, None, None, None, None


LEAKY_SLOPE = 0.01     # nn.LeakyReLU's default; csrc/igemm.cuh kLeakySlope


def conv_apply(state, x0, x1, weight, bias, relu=False, mode=0, out_hw=None, link=None, side=False):
    relu = 2 if relu == "leaky" else int(relu)          # False / True / "leaky" -> the kernels' activation id
    x0 = nhwc(x0, BF16)
    if x1 is not None:
        x1 = nhwc(x1, BF16)
        if x1.shape[0] != x0.shape[0] or x1.shape[2:] != x0.shape[2:]:
            raise ValueError(f"concat sources disagree: {tuple(x0.shape)} vs {tuple(x1.shape)}")
    if x0.shape[1] != state.c0 or (x1 is not None and x1.shape[1] != state.c1):
        raise ValueError(f"expected {state.c0}+{state.c1} input channels, got {x0.shape[1]}")
    return _ConvFn.apply(x0, x1, weight, bias, state, relu, mode, out_hw, link, side)


# ====================================================================== GDN / IGDN
PEDESTAL = 2.0 ** -36   # reparam_offset**2, compressai/ops/parametrizers.py:47-53


class GDNState:
    """Per-layer GDN state.  Stand-alone modules reparametrise beta / gamma on every call; inside a
    model's ParamArena the effective values (`beta_eff` fp32, `gamma_bf` bf16) and the buffers that
    receive their gradients (`g_beta_eff`, `g_gamma_eff`) are persistent views filled for ALL GDN
    layers by one batched launch per direction (ParamArena.prepare / finish_backward)."""

    def __init__(self, C, inverse, beta_bound, gamma_bound):
        self.C, self.inverse = C, bool(inverse)
        self.beta_bound, self.gamma_bound = float(beta_bound), float(gamma_bound)
        self.grad_ready = None
        self.arena = None
        self.link = None           # GDNLink of the latest forward pass (conv_gdn_apply)
        self.beta_eff = self.gamma_bf = self.g_beta_eff = self.g_gamma_eff = None


def wait_gdn_gate(arena):
    """Pipelined train step: the reparametrised beta / gamma of this step are produced on the optimiser stream behind the
    first slice's update (train.TrainStep._apply_pending); the first GDN launch of the forward pass waits for them (every
    later stream of the forward pass forks from this one behind that point)."""
    if arena is not None and arena.gates:
        ev = arena.gates.pop("gdn", None)
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)


class _GDNFn(torch.autograd.Function):
    """compressai/layers/gdn.py:77-92 with the NonNegativeParametrizer of beta / gamma
    (ops/parametrizers.py:60-63) and LowerBound's custom gradient (ops/bound_ops.py:40-42)."""

    @staticmethod
    def forward(ctx, x, beta, gamma, state):
        B, C, H, W = x.shape
        P = B * H * W
        dev, st = x.device, stream()
        arena = state.arena
        if arena is not None:
            arena.ensure_prepared(beta, gamma)
            beta_eff, gamma_bf = state.beta_eff, state.gamma_bf
        else:
            beta_eff = torch.empty(C, dtype=F32, device=dev)
            gamma_bf = torch.empty((C, C), dtype=BF16, device=dev)
            L.reparam_forward(ptr(beta), C, state.beta_bound, PEDESTAL, ptr(beta_eff), None, st)
            L.reparam_forward(ptr(gamma), C * C, state.gamma_bound, PEDESTAL, None, ptr(gamma_bf), st)
        wait_gdn_gate(arena)
        train = any(ctx.needs_input_grad[:3])
        x2 = torch.empty_like(x) if train else None      # inference: x^2 never leaves the GEMM's producer warps
        y = torch.empty_like(x)
        rn = torch.empty_like(x)
        L.gdn_forward(ptr(x), ptr(gamma_bf), ptr(beta_eff), int(state.inverse), P, C, ptr(x2), ptr(y), ptr(rn), st)
        ctx.state, ctx.beta, ctx.gamma = state, beta, gamma
        if train:
            ctx.save_for_backward(x, rn, gamma_bf, x2)   # x^2 is kept: 180 GB of HBM, no recompute pass in backward
        return y

    @staticmethod
    def backward(ctx, gy):
        state = ctx.state
        x, rn, gamma_bf, x2 = ctx.saved_tensors
        B, C, H, W = x.shape
        P = B * H * W
        dev, st = x.device, stream()
        gy = nhwc(gy, BF16)
        t = torch.empty_like(x)
        gx = torch.empty_like(x)
        arena = state.arena
        if arena is not None:
            g_gamma_eff, g_beta_eff = state.g_gamma_eff, state.g_beta_eff
        else:
            g_gamma_eff = torch.zeros((C, C), dtype=F32, device=dev)
            g_beta_eff = torch.zeros(C, dtype=F32, device=dev)
        ws = SideStreams.next_wgrad() if arena is not None else None
        L.gdn_backward(ptr(gy), ptr(x), ptr(rn), ptr(gamma_bf), int(state.inverse), P, C, ptr(x2), 1, ptr(t), ptr(gx),
                       None if ws is not None else ptr(g_gamma_eff), None if ws is not None else ptr(g_beta_eff), st)
        if ws is not None:                       # gamma's gradient GEMM is off the critical path
            ev = torch.cuda.Event()
            ev.record()
            t.record_stream(ws)
            x2.record_stream(ws)
            with torch.cuda.stream(ws):
                ws.wait_event(ev)
                L.gdn_wgrad(ptr(t), ptr(x2), P, C, ptr(g_gamma_eff), ptr(g_beta_eff), stream())   # gamma and beta gradients
        beta, gamma = ctx.beta, ctx.gamma
        if arena is not None:
            arena.queue_finish_backward()       # one batched reparam-backward at the end of this backward pass
        else:
            gb, gg = grad_view(beta), grad_view(gamma)
            L.reparam_backward(ptr(g_beta_eff), ptr(beta), C, state.beta_bound, ptr(gb), st)
            L.reparam_backward(ptr(g_gamma_eff), ptr(gamma), C * C, state.gamma_bound, ptr(gg), st)
            if state.grad_ready is not None:
                state.grad_ready([beta, gamma])
        return (gx if ctx.needs_input_grad[0] else None), None, None, None


def gdn_apply(state, x, beta, gamma):
    return _GDNFn.apply(nhwc(x, BF16), beta, gamma, state)


class _ConvGDNFn(torch.autograd.Function):
    """conv / deconv followed by GDN / IGDN as ONE node (models/utils.py:128-146 + layers/gdn.py:77-92).

    Forward: when the conv's launch uses whole-K 256-wide tiles anyway (lmsfc_conv_gdn_fusable) the GDN's channel
    contraction runs in the conv kernel's epilogue on the accumulator tile (lmsfc_conv_gdn_forward); otherwise the two
    kernels run back to back.  Backward: the GDN part is skipped when the consumer of y already ran it inside its
    data-gradient kernel (GDNLink.fused_done: the incoming gradient then IS the gradient w.r.t. the conv output); the
    conv part hands ITS data gradient to the GDN of the layer before (`in_link`) the same way.
    Returns y; the GDNLink for the consumers of y is left in `gdn_state.link`."""

    @staticmethod
    def forward(ctx, x0, x1, weight, bias, beta, gamma, cstate, gstate, out_hw, in_link):
        B, _, H, W = x0.shape
        Ho, Wo = cstate.out_hw(H, W, out_hw)
        C = cstate.cout
        cstate.refresh(weight)
        d = cstate.desc(B, H, W, Ho, Wo)
        dev, st = x0.device, stream()
        arena = gstate.arena
        if arena is not None:
            arena.ensure_prepared(beta, gamma)
            beta_eff, gamma_bf = gstate.beta_eff, gstate.gamma_bf
        else:
            beta_eff = torch.empty(C, dtype=F32, device=dev)
            gamma_bf = torch.empty((C, C), dtype=BF16, device=dev)
            L.reparam_forward(ptr(beta), C, gstate.beta_bound, PEDESTAL, ptr(beta_eff), None, st)
            L.reparam_forward(ptr(gamma), C * C, gstate.gamma_bound, PEDESTAL, None, ptr(gamma_bf), st)
        train = any(ctx.needs_input_grad[:6])
        y = empty_nhwc(B, C, Ho, Wo, BF16, dev)
        x = rn = x2 = None
        if train:
            x, rn, x2 = (empty_nhwc(B, C, Ho, Wo, BF16, dev) for _ in range(3))
        if C == 256 and L.gdn_fusable(d, inference=not train):
            wait_gdn_gate(arena)
            L.conv_gdn_forward(d, ptr(x0), ptr(x1), ptr(cstate.w_direct), ptr(bias), ptr(gamma_bf), ptr(beta_eff),
                               int(gstate.inverse), ptr(x), ptr(x2), ptr(y), ptr(rn), st)
        else:
            if x is None:
                x = empty_nhwc(B, C, Ho, Wo, BF16, dev)
            L.conv_forward(d, ptr(x0), ptr(x1), ptr(cstate.w_direct), ptr(bias), 0, ptr(x), None, None, st)
            wait_gdn_gate(arena)
            L.gdn_forward(ptr(x), ptr(gamma_bf), ptr(beta_eff), int(gstate.inverse), B * Ho * Wo, C, ptr(x2), ptr(y), ptr(rn), st)
        ctx.cstate, ctx.gstate, ctx.shape = cstate, gstate, (B, H, W, Ho, Wo)
        ctx.weight, ctx.bias, ctx.beta, ctx.gamma = weight, bias, beta, gamma
        ctx.need_dx = ctx.needs_input_grad[0]
        ctx.in_link = in_link
        ctx.in_role = in_link.register_consumer(cstate, d, False) if (in_link is not None and ctx.need_dx) else None
        link = GDNLink(x, rn, gamma_bf, gstate.inverse) if train else None
        ctx.link = gstate.link = link
        if train:
            ctx.save_for_backward(x0, x1, x, rn, gamma_bf, x2)
        return y

    @staticmethod
    def backward(ctx, gy):
        cstate, gstate, link = ctx.cstate, ctx.gstate, ctx.link
        x0, x1, x, rn, gamma_bf, x2 = ctx.saved_tensors
        B, H, W, Ho, Wo = ctx.shape
        C, P = cstate.cout, B * Ho * Wo
        dev, st = x.device, stream()
        gy = nhwc(gy, BF16)
        arena = gstate.arena
        if arena is not None:
            g_gamma_eff, g_beta_eff = gstate.g_gamma_eff, gstate.g_beta_eff
        else:
            g_gamma_eff = torch.zeros((C, C), dtype=F32, device=dev)
            g_beta_eff = torch.zeros(C, dtype=F32, device=dev)
        ws = SideStreams.next_wgrad() if arena is not None else None
        if link.fused_done:                      # the consumer's dgrad kernel ran the contraction: gy is d loss / d x
            gx, t = gy, link.t
        else:
            t = torch.empty_like(x)
            gx = torch.empty_like(x)
            L.gdn_backward(ptr(gy), ptr(x), ptr(rn), ptr(gamma_bf), int(gstate.inverse), P, C, ptr(x2), 1, ptr(t), ptr(gx),
                           None, None, st)
        if ws is not None:                       # gamma's / beta's gradient GEMM is off the critical path
            ev = torch.cuda.Event()
            ev.record()
            t.record_stream(ws)
            x2.record_stream(ws)
            with torch.cuda.stream(ws):
                ws.wait_event(ev)
                L.gdn_wgrad(ptr(t), ptr(x2), P, C, ptr(g_gamma_eff), ptr(g_beta_eff), stream())
        else:
            L.gdn_wgrad(ptr(t), ptr(x2), P, C, ptr(g_gamma_eff), ptr(g_beta_eff), st)
        beta, gamma = ctx.beta, ctx.gamma
        if arena is not None:
            arena.queue_finish_backward()        # one batched reparam-backward at the end of this backward pass
        else:
            gb, gg = grad_view(beta), grad_view(gamma)
            L.reparam_backward(ptr(g_beta_eff), ptr(beta), C, gstate.beta_bound, ptr(gb), st)
            L.reparam_backward(ptr(g_gamma_eff), ptr(gamma), C * C, gstate.gamma_bound, ptr(gg), st)
            if gstate.grad_ready is not None:
                gstate.grad_ready([beta, gamma])
        link.t = link.addend = None              # drop the per-pass tensors
        gin = _conv_backward(cstate, ctx.shape, x0, x1, gx, ctx.weight, ctx.bias, ctx.need_dx, ctx.in_link, ctx.in_role)
        return gin, None, None, None, None, None, None, None, None, None


def conv_gdn_apply(cstate, gstate, x0, x1, weight, bias, beta, gamma, out_hw=None, in_link=None):
    """-> (y, GDNLink or None)."""
    x0 = nhwc(x0, BF16)
    if x1 is not None:
        x1 = nhwc(x1, BF16)
    if x0.shape[1] != cstate.c0 or (x1 is not None and x1.shape[1] != cstate.c1) or cstate.cout != gstate.C:
        raise ValueError(f"conv_gdn: channels {x0.shape[1]} / {cstate.c0}+{cstate.c1} -> {cstate.cout} / GDN({gstate.C})")
    y = _ConvGDNFn.apply(x0, x1, weight, bias, beta, gamma, cstate, gstate, out_hw, in_link)
    return y, gstate.link


# ====================================================================== entropy models
MODES = {"noise": 0, "dequantize": 1}


def _mode_id(mode):
    if mode not in MODES:
        raise ValueError(f'Invalid quantization mode: "{mode}"')   # entropy_models.py:158-159
    return MODES[mode]


class RNG:
    """Counter-based Philox stream for the quantisation noise: {seed, step} lives on the device so
    a captured CUDA graph draws fresh noise at every replay; `offset` separates draws in a step."""

    def __init__(self, device, seed=0):
        self.state = torch.tensor([seed, 0], dtype=torch.int64, device=device)
        self.seed = int(seed)

    def advance(self):
        L.rng_advance(ptr(self.state), stream())


_SUM_LN_ATTR = "_lmsfc_sum_ln"        # [1] fp32: sum(ln likelihood), reduced inside the likelihood kernel
_BF16_ATTR = "_lmsfc_bf16"            # bf16 copy of the quantised latent (operand of the next conv)
_RATE_ATTR = "_lmsfc_rate_hook"       # RateHook shared between the likelihood op and the RD loss
_CARRIED = (_SUM_LN_ATTR, _BF16_ATTR, _RATE_ATTR)


def carry_attrs(src, dst):
    """Views created by reshape() drop Python attributes: copy the side-band ones over."""
    for a in _CARRIED:
        v = getattr(src, a, None)
        if v is not None:
            setattr(dst, a, v)
    return dst


class RateHook:
    """Lets RateDistortionLoss hand d loss / d likelihood = coef / L to the likelihood op's own
    backward kernel (no reciprocal / scale passes, no gradient tensor)."""
    __slots__ = ("coef",)

    def __init__(self):
        self.coef = None


class _GCFn(torch.autograd.Function):
    """GaussianConditional.forward, entropy_models.py:650-682.
    Returns (y_hat fp32, likelihood fp32, y_hat bf16)."""

    @staticmethod
    def forward(ctx, y, scales, means, noise, mode, scale_bound, lik_bound, rng, offset):
        n = y.numel()
        y_hat = torch.empty_like(y)
        y_hat_b = torch.empty_like(y, dtype=BF16)
        lik = torch.empty_like(y)
        sum_ln = torch.zeros(1, dtype=F32, device=y.device)
        L.gc_forward(ptr(y), ptr(scales), ptr(means), ptr(noise), n, mode, scale_bound, lik_bound,
                     rng.seed if rng is not None else 0, offset, ptr(rng.state) if rng is not None else None,
                     ptr(y_hat), ptr(y_hat_b), ptr(lik), ptr(sum_ln), stream())
        ctx.save_for_backward(y_hat, scales, means)
        ctx.cfg = (mode, scale_bound, lik_bound)
        ctx.hook = RateHook()
        ctx.set_materialize_grads(False)      # unused outputs arrive as None, not as zero tensors
        setattr(lik, _SUM_LN_ATTR, sum_ln)
        setattr(lik, _RATE_ATTR, ctx.hook)
        return y_hat, lik, y_hat_b

    @staticmethod
    def backward(ctx, g_yhat, g_lik, g_yhat_b):
        y_hat, scales, means = ctx.saved_tensors
        mode, scale_bound, lik_bound = ctx.cfg
        n = y_hat.numel()
        g_y = torch.empty_like(y_hat)
        g_s = torch.empty_like(y_hat)
        g_m = torch.empty_like(y_hat) if means is not None else None
        rate_coef, ctx.hook.coef = (ctx.hook.coef or 0.0), None
        if g_lik is None and rate_coef == 0.0:
            g_lik = torch.zeros_like(y_hat)
        if g_lik is not None:
            g_lik = g_lik.to(dtype=F32).contiguous(memory_format=CL)
        if g_yhat is not None and g_yhat_b is not None:
            g_yhat, g_yhat_b = g_yhat.to(F32) + g_yhat_b.to(F32), None
        if g_yhat is not None:
            g_yhat = g_yhat.to(dtype=F32).contiguous(memory_format=CL)
        if g_yhat_b is not None:
            g_yhat_b = g_yhat_b.to(dtype=BF16).contiguous(memory_format=CL)
        L.gc_backward(ptr(y_hat), ptr(scales), ptr(means), ptr(g_lik), ptr(g_yhat), ptr(g_yhat_b), n, mode, scale_bound,
                      lik_bound, rate_coef, ptr(g_y), ptr(g_s), ptr(g_m), stream())
        return g_y, g_s, g_m, None, None, None, None, None, None


def _al16(t):
    """The vectorised entropy kernels use 16-byte accesses: re-home a view that starts at an odd offset."""
    return t if (t is None or t.data_ptr() % 16 == 0) else t.clone(memory_format=CL)


def gc_apply(y, scales, means, noise, mode, scale_bound, lik_bound, rng=None, offset=1):
    y = _al16(nhwc(y, F32))
    scales = _al16(nhwc(scales, F32))
    if means is not None:
        means = _al16(nhwc(means, F32))
    if noise is not None:
        noise = _al16(nhwc(noise, F32))
    out, lik, out_b = _GCFn.apply(y, scales, means, noise, _mode_id(mode), float(scale_bound), float(lik_bound), rng, offset)
    setattr(out, _BF16_ATTR, out_b)
    return out, lik


class _QuantNoiseFn(torch.autograd.Function):
    """y_hat = y + U(-.5, .5) (entropy_models.py:161-165) as its own node: in noise mode the quantised
    latent does not depend on the hyperprior, so the synthesis transform can start while the hyper path
    (and later the likelihood) runs on another stream.  Returns (y_hat fp32, y_hat bf16)."""

    @staticmethod
    def forward(ctx, y, noise, rng, offset):
        y_hat = torch.empty_like(y)
        y_hat_b = torch.empty_like(y, dtype=BF16)
        L.gc_forward(ptr(y), None, None, ptr(noise), y.numel(), 0, 0.0, 0.0, rng.seed if rng is not None else 0, offset,
                     ptr(rng.state) if rng is not None else None, ptr(y_hat), ptr(y_hat_b), None, None, stream())
        ctx.set_materialize_grads(False)
        return y_hat, y_hat_b

    @staticmethod
    def backward(ctx, g, g_b):
        if g is None:
            out = None if g_b is None else g_b.to(F32)
        else:
            out = g if g_b is None else torch.add(g, g_b)      # fp32 + bf16 -> fp32, one pass
        return out, None, None, None


def quant_noise_apply(y, noise=None, rng=None, offset=1):
    y = _al16(nhwc(y, F32))
    if noise is not None:
        noise = _al16(nhwc(noise, F32))
    return _QuantNoiseFn.apply(y, noise, rng, offset)


def gc_likelihood_apply(y_hat, scales, scale_bound, lik_bound, means=None):
    """Gaussian likelihood of an already quantised latent (mode "given"); same node type as gc_apply."""
    if means is not None:
        means = _al16(nhwc(means, F32))
    _, lik, _ = _GCFn.apply(_al16(nhwc(y_hat, F32)), _al16(nhwc(scales, F32)), means, None, 2, float(scale_bound), float(lik_bound), None, 0)
    return lik


class _EBFn(torch.autograd.Function):
    """EntropyBottleneck.forward, entropy_models.py:436-516, on the NHWC view [P][C].
    Returns (y_hat fp32, likelihood fp32, y_hat bf16).  `packed_grad` (optional, [C,58] fp32 view of
    the gradient arena): the backward kernel accumulates the parameter gradients straight into it."""

    @staticmethod
    def forward(ctx, x, packed, medians, noise, mode, lik_bound, rng, offset, packed_grad):
        B, C, H, W = x.shape
        P = B * H * W
        y_hat = torch.empty_like(x)
        y_hat_b = torch.empty_like(x, dtype=BF16)
        lik = torch.empty_like(x)
        sum_ln = torch.zeros(1, dtype=F32, device=x.device)
        L.eb_forward(ptr(x), ptr(packed), ptr(medians), ptr(noise), P, C, mode, lik_bound,
                     rng.seed if rng is not None else 0, offset, ptr(rng.state) if rng is not None else None,
                     ptr(y_hat), ptr(y_hat_b), ptr(lik), ptr(sum_ln), stream())
        ctx.save_for_backward(y_hat, packed)
        ctx.cfg = (mode, lik_bound, P, C)
        ctx.packed_grad = packed_grad
        ctx.hook = RateHook()
        ctx.set_materialize_grads(False)      # unused outputs arrive as None, not as zero tensors
        setattr(lik, _SUM_LN_ATTR, sum_ln)
        setattr(lik, _RATE_ATTR, ctx.hook)
        return y_hat, lik, y_hat_b

    @staticmethod
    def backward(ctx, g_yhat, g_lik, g_yhat_b):
        y_hat, packed = ctx.saved_tensors
        mode, lik_bound, P, C = ctx.cfg
        g_x = torch.empty_like(y_hat)
        g_p = ctx.packed_grad if ctx.packed_grad is not None else torch.zeros_like(packed)
        rate_coef, ctx.hook.coef = (ctx.hook.coef or 0.0), None
        if g_lik is None and rate_coef == 0.0:
            g_lik = torch.zeros_like(y_hat)
        if g_lik is not None:
            g_lik = g_lik.to(dtype=F32).contiguous(memory_format=CL)
        if g_yhat is not None and g_yhat_b is not None:
            g_yhat, g_yhat_b = g_yhat.to(F32) + g_yhat_b.to(F32), None
        if g_yhat is not None:
            g_yhat = g_yhat.to(dtype=F32).contiguous(memory_format=CL)
        if g_yhat_b is not None:
            g_yhat_b = g_yhat_b.to(dtype=BF16).contiguous(memory_format=CL)
        L.eb_backward(ptr(y_hat), ptr(packed), ptr(g_lik), ptr(g_yhat), ptr(g_yhat_b), P, C, mode, lik_bound, rate_coef,
                      ptr(g_x), ptr(g_p), stream())
        return g_x, (None if ctx.packed_grad is not None else g_p), None, None, None, None, None, None, None


def eb_apply(x, packed, medians, noise, mode, lik_bound, rng=None, offset=0, packed_grad=None):
    x = nhwc(x, F32)
    if noise is not None:
        noise = nhwc(noise, F32)
    out, lik, out_b = _EBFn.apply(x, packed.contiguous(), medians, noise, _mode_id(mode), float(lik_bound), rng, offset,
                                  packed_grad)
    setattr(out, _BF16_ATTR, out_b)
    return out, lik


class _EBAuxFn(torch.autograd.Function):
    """EntropyBottleneck.loss, entropy_models.py:431-434 (gradient to `quantiles` only)."""

    @staticmethod
    def forward(ctx, quantiles, packed, target):
        C = quantiles.shape[0]
        q = quantiles.detach().reshape(C, 3).contiguous()
        loss = torch.zeros(1, dtype=F32, device=q.device)
        gq = torch.empty_like(q)
        L.eb_aux_loss(ptr(packed), ptr(q), ptr(target), C, ptr(loss), ptr(gq), stream())
        ctx.save_for_backward(gq)
        ctx.qshape = quantiles.shape
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        (gq,) = ctx.saved_tensors
        return (gq * g).reshape(ctx.qshape), None, None


def eb_aux_apply(quantiles, packed, target):
    return _EBAuxFn.apply(quantiles, packed.detach().contiguous(), target)


# ====================================================================== RD loss
_CONST_CACHE = {}


def _const_tensor(values, device):
    """Small fp32 constant vectors, uploaded once per (values, device): the training step must not
    issue host->device copies of its own (it is captured into a CUDA graph)."""
    key = (values, str(device))
    t = _CONST_CACHE.get(key)
    if t is None:
        t = torch.tensor(values, dtype=F32, device=device)
        _CONST_CACHE[key] = t
    return t


class _RDLossFn(torch.autograd.Function):
    """RateDistortionLoss, compressai/losses/rate_distortion.py:47-59, for a pyramid of targets:
    bpp = sum_k sum ln L_k / (-ln2 * num_pixels); D = sum_l w_l * mean((xhat_l - x_l)^2);
    loss = lmbda_scale * D + bpp.  One pass per level produces both the squared error and the
    gradient d loss / d xhat_l (the loss is assumed to be the root of backward)."""

    @staticmethod
    def forward(ctx, lmbda_scale, num_pixels, weights, n_lev, n_lik, extra, *tensors):
        xh = tensors[:n_lev]
        xt = tensors[n_lev:2 * n_lev]
        liks = tensors[2 * n_lev:2 * n_lev + n_lik]
        dev, st = xh[0].device, stream()
        ctx.set_materialize_grads(False)          # bpp / mse carry no gradient: no zero tensors made for them
        rate_c = 1.0 / (-math.log(2) * num_pixels)
        import ctypes
        grads = [torch.empty_like(xh[i]) if ctx.needs_input_grad[6 + i] else None for i in range(n_lev)]
        ns = [xh[i].numel() for i in range(n_lev)]
        VP, I64, F32A = ctypes.c_void_p * n_lev, ctypes.c_int64 * n_lev, ctypes.c_float * n_lev
        gcoef = F32A(*[lmbda_scale * weights[i] * 2.0 / ns[i] for i in range(n_lev)])
        sums = [getattr(lk, _SUM_LN_ATTR, None) for lk in liks]
        if all(s_ is not None for s_ in sums) and n_lik <= 8:
            # the whole loss in ONE launch behind the last reconstruction: squared errors + gradients of all levels,
            # fixed-order reduction, bpp / D / loss (+ the aux loss riding along) by the last block (lmsfc_rd_loss)
            out4 = torch.empty(4, dtype=F32, device=dev)
            RP = ctypes.c_void_p * max(n_lik, 1)
            L.rd_loss(n_lev, VP(*[ptr(t) for t in xh]), VP(*[ptr(t) for t in xt]), I64(*ns), gcoef,
                      VP(*[ptr(g) for g in grads]), n_lik, RP(*[ptr(s_) for s_ in sums]), rate_c,
                      F32A(*[weights[i] / ns[i] for i in range(n_lev)]), lmbda_scale, ptr(extra), ptr(out4), st)
            out = (out4[0], out4[1], out4[2])
            scalars4 = out4.detach()
        else:
            accum = torch.zeros(n_lik + n_lev, dtype=F32, device=dev)
            coef = [rate_c] * n_lik
            for i, lk in enumerate(liks):
                accum[i:i + 1] = sums[i] if sums[i] is not None else torch.log(lk.float()).sum()
            L.mse_forward_backward_multi(n_lev, VP(*[ptr(t) for t in xh]), VP(*[ptr(t) for t in xt]), I64(*ns), gcoef,
                                         ptr(accum[n_lik:]), VP(*[ptr(g) for g in grads]), st)   # all levels, one launch
            coef += [weights[i] / ns[i] for i in range(n_lev)]
            coef_t = _const_tensor(tuple(coef), dev)
            o3 = torch.empty(3, dtype=F32, device=dev)
            L.loss_finalize(ptr(accum), ptr(coef_t), n_lik, n_lev, lmbda_scale, ptr(o3), st)
            out = (o3[2], o3[0], o3[1])
            scalars4 = None
        # likelihoods produced by the fused entropy kernels take d loss / d L = rate_c / L inside their
        # own backward (RateHook); any other likelihood tensor gets an explicit gradient
        ctx.fused = []
        for lk in liks:
            hook = getattr(lk, _RATE_ATTR, None)
            ctx.fused.append(hook is not None)
            if hook is not None:
                hook.coef = rate_c
        ctx.grads, ctx.rate_c, ctx.n_lev = grads, rate_c, n_lev
        ctx.save_for_backward(*[lk for lk, f in zip(liks, ctx.fused) if not f])
        if scalars4 is not None:
            ctx.mark_non_differentiable(scalars4)
        return (*out, scalars4)                  # the fourth output carries no gradient (None on the fallback path)

    @staticmethod
    def backward(ctx, g_loss, g_bpp, g_mse, g_scalars=None):
        plain = iter(ctx.saved_tensors)
        g_liks = [None if f else ctx.rate_c / next(plain) for f in ctx.fused]
        return (None, None, None, None, None, None, *ctx.grads, *([None] * ctx.n_lev), *g_liks)


def rd_loss_apply(x_hats, targets, liks, lmbda_scale, num_pixels, weights, extra=None, with_scalars=False):
    """-> (loss, bpp, mse) [, scalars4]: scalars4 = the device vector {loss, bpp, mse, *extra} the loss kernel wrote (None
    on the fallback path); `extra`: a device scalar to ride along (the aux loss of a train step)."""
    xh = [nhwc(t, BF16) for t in x_hats]
    xt = [nhwc(t, BF16) for t in targets]
    for a, b in zip(xh, xt):
        if a.shape != b.shape:
            raise ValueError(f"x_hat {tuple(a.shape)} vs target {tuple(b.shape)}")
    if extra is not None:
        extra = extra.detach().reshape(-1).to(F32)
    loss, bpp, mse, s4 = _RDLossFn.apply(float(lmbda_scale), float(num_pixels), [float(w) for w in weights], len(xh), len(liks),
                                         extra, *xh, *xt, *liks)
    return (loss, bpp, mse, s4) if with_scalars else (loss, bpp, mse)
