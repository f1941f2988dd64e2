"""The L-MSFC codec model (topology frozen in spec.py) on the B200 kernels.

Contract kept from compressai.models.base.CompressionModel (models/base.py:62-144) and
models/google.py:292-303: `forward(x) -> {"x_hat": ..., "likelihoods": {...}}`, `aux_loss()`,
`update(scale_table=None, force=False)`, `load_state_dict` resizing the CDF buffers first.
Here `x` and `x_hat` are dicts of the pyramid levels (p2..p5, [B,F,H,W]).

ParamArena packs every fp32 parameter (and its gradient) into one flat allocation so that
clip+Adam is one launch, the bf16 GEMM operands of all conv weights are one contiguous block
written by the same launch (each is read K-major or MN-major by the kernels: no transposed copies), and gradient all-reduce buckets are plain slices.
"""
import math

import torch
import torch.nn as nn

from . import _lib as L
from . import functional as Fn
from ._lib import ptr, stream
from .entropy_models import EntropyBottleneck, GaussianConditional
from .layers import GDN, Conv, Deconv, _ConvBase
from .spec import DEFAULT_SPEC, LMSFCSpec

SCALES_MIN, SCALES_MAX, SCALES_LEVELS = 0.11, 256, 64


def get_scale_table(min=SCALES_MIN, max=SCALES_MAX, levels=SCALES_LEVELS):
    """compressai/models/base.py:57-59."""
    return torch.exp(torch.linspace(math.log(min), math.log(max), levels))


class LMSFC(nn.Module):
    def __init__(self, spec: LMSFCSpec = DEFAULT_SPEC):
        super().__init__()
        self.spec = spec
        N, M, ch = spec.N, spec.M, spec.channels
        if N % 256 or M % 256 or any(c % 256 for c in ch):
            # csrc: the weight-gradient tiles are 256 x 256 channels (wgrad.cu), GDN's contraction needs C % 256 == 0
            raise ValueError(f"LMSFC on the B200 kernels: N, M and the level channels must be multiples of 256, got {N}, {M}, {ch}")
        n = len(spec.levels)
        self._tags = [spec.tag(i) for i in range(n)]
        g_a = nn.Module()
        for i, t in enumerate(self._tags):        # conv<t> (+ gdn<t>) per level; the last one produces y
            last = i == n - 1
            cin = ch[i] + (N if i else 0)
            setattr(g_a, f"conv{t}", Conv(cin, M if last else N, concat_channels=ch[i] if i else 0))
            if not last:
                setattr(g_a, f"gdn{t}", GDN(N))
        self.g_a = g_a
        # keys "0","2","4" mirror the nn.Sequential(conv, ReLU, conv, ReLU, conv) of google.py:267-282
        self.h_a = nn.ModuleDict({"0": Conv(M, N, 3, 1), "2": Conv(N, N), "4": Conv(N, N)})
        if spec.mean_scale:       # models/google.py:377-391 (LeakyReLU between the layers, no final activation)
            if spec.mid % 256:
                raise ValueError(f"LMSFC on the B200 kernels: hyper_mid must be a multiple of 256, got {spec.mid}")
            self.h_s = nn.ModuleDict({"0": Deconv(N, M), "2": Deconv(M, spec.mid), "4": Conv(spec.mid, 2 * M, 3, 1)})
        else:
            self.h_s = nn.ModuleDict({"0": Deconv(N, N), "2": Deconv(N, N), "4": Conv(N, M, 3, 1)})
        self._act = "leaky" if spec.mean_scale else True
        self.entropy_bottleneck = EntropyBottleneck(N)
        self.gaussian_conditional = GaussianConditional(None, scale_bound=spec.scale_bound,
                                                        likelihood_bound=spec.likelihood_bound)
        g_s = nn.Module()
        for i in range(n - 1, 0, -1):             # up<t> + igdn<t> + head<t> from the coarsest level down; up<t0> ends
            t = self._tags[i]
            setattr(g_s, f"up{t}", Deconv(M if i == n - 1 else N, N))
            setattr(g_s, f"igdn{t}", GDN(N, inverse=True))
            setattr(g_s, f"head{t}", Conv(N, ch[i], 3, 1))
        setattr(g_s, f"up{self._tags[0]}", Deconv(N, ch[0]))
        self.g_s = g_s
        self.arena = None
        self.entropy_bottleneck.rng_auto_advance = False
        self.gaussian_conditional.rng_auto_advance = False
        self.GATES = spec.gate_names()            # first parameter of every optimiser slice, in arena (= forward) order

    def _share_rng(self, device):
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        gc.rng = eb._rng(device)

    # ------------------------------------------------------------------ forward
    def forward(self, feats, noise_z=None, noise_y=None):
        """feats: {"p2": [B,F,H,W], ...}.  noise_z / noise_y (optional, shaped like z / y) inject
        the U(-.5,.5) draws for parity tests; RNG order is EB then GC (google.py:295-297)."""
        lv = self.spec.levels
        ps = tuple(Fn.nhwc(feats[k], Fn.BF16) for k in lv)
        self._share_rng(ps[0].device)
        level_hw = [p.shape[2:] for p in ps]
        hs = Fn.SideStreams.hyper if self.training else None
        if hs is None:
            y, z, z2_hw = self.analysis(ps)
            z_hat, z_lik = self.entropy_bottleneck(z, noise=noise_z)
            if self.training:
                scales, means = self.hyper_synthesis(getattr(z_hat, Fn._BF16_ATTR), z2_hw, y.shape[2:])
            else:                  # eval: the bits compress() / decompress() compute (batch-invariant tiling)
                with L.batch_invariant():
                    scales, means = self.hyper_synthesis(getattr(z_hat, Fn._BF16_ATTR), z2_hw, y.shape[2:])
            y_hat, y_lik = self.gaussian_conditional(y, scales, means=means, noise=noise_y)
            x_hat = self.synthesis(getattr(y_hat, Fn._BF16_ATTR), level_hw)
        else:
            # noise mode: y_hat = y + U needs no hyperprior -> g_s starts right away on this stream while
            # h_a -> EntropyBottleneck -> h_s -> Gaussian likelihood run on the hyper stream; autograd replays
            # each node on the stream it was recorded on, so the two backward chains overlap as well.
            eb, gc = self.entropy_bottleneck, self.gaussian_conditional
            y, y_abs = self.encode_y(ps)
            y_hat, y_hat_b = Fn.quant_noise_apply(y, noise_y, gc._rng(y.device) if noise_y is None else None, gc.rng_offset)
            main = torch.cuda.current_stream()
            ev = torch.cuda.Event()
            ev.record(main)
            for t in (y_abs, y_hat):
                t.record_stream(hs)
            with torch.cuda.stream(hs):
                hs.wait_event(ev)
                z, z2_hw = self.hyper_analysis(y_abs)
                z_hat, z_lik = eb(z, noise=noise_z)
                scales, means = self.hyper_synthesis(getattr(z_hat, Fn._BF16_ATTR), z2_hw, y.shape[2:])
                y_lik = Fn.gc_likelihood_apply(y_hat, scales, gc.scale_bound_value,
                                               gc.likelihood_bound if gc.use_likelihood_bound else 0.0, means=means)
            x_hat = self.synthesis(y_hat_b, level_hw)
            main.wait_stream(hs)
            for t in (z, z_hat, z_lik, scales, means, y_lik, getattr(z_lik, Fn._SUM_LN_ATTR), getattr(y_lik, Fn._SUM_LN_ATTR)):
                if t is not None:
                    t.record_stream(main)
        if self.training and noise_z is None and noise_y is None:
            self.entropy_bottleneck.rng.advance()      # ONE shared counter bump per step (EB and GC share the RNG)
        return {"x_hat": dict(zip(lv, x_hat)),
                "likelihoods": {"y": y_lik, "z": z_lik},
                "latents": {"y": y, "z": z, "y_hat": y_hat, "z_hat": z_hat, "scales": scales, "means": means},
                "inputs_nhwc": dict(zip(lv, ps))}

    def encode_y(self, ps):
        """g_a: (pyramid levels, bf16 NHWC) -> y fp32 and the input of h_a in bf16: |y| (models/google.py:294), under the
        mean-scale hyperprior y itself (:395)."""
        a, tags = self.g_a, self._tags
        # conv + GDN are one node each; the GDNLink lets the NEXT layer's data-gradient kernel run this GDN's backward
        e, link = getattr(a, f"conv{tags[0]}").forward_gdn(getattr(a, f"gdn{tags[0]}"), ps[0])
        for i in range(1, len(tags) - 1):
            self._gate(f"g_a.conv{tags[i]}.weight")
            e, link = getattr(a, f"conv{tags[i]}").forward_gdn(getattr(a, f"gdn{tags[i]}"), e, ps[i], link=link)
        self._gate(f"g_a.conv{tags[-1]}.weight")
        return getattr(a, f"conv{tags[-1]}")(e, ps[-1], latent_head="plain" if self.spec.mean_scale else True, link=link)

    def _gate(self, name):
        """Pipelined train step: wait (on the current stream) for the optimiser slice that starts at this parameter."""
        if self.arena is not None and self.arena.gates:
            ev = self.arena.gates.get(name)
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)

    def hyper_analysis(self, y_abs):
        """h_a: |y| -> z fp32 and the spatial size of its middle activation."""
        self._gate("h_a.0.weight")
        z1 = self.h_a["0"](y_abs, relu=self._act)
        z2 = self.h_a["2"](z1, relu=self._act)
        z, _ = self.h_a["4"](z2, latent_head=True)
        return z, z2.shape[2:]

    def analysis(self, ps):
        """g_a + h_a: (p2..p5 bf16 NHWC) -> y fp32, z fp32, spatial size of h_a's middle activation."""
        y, y_abs = self.encode_y(ps)
        z, z2_hw = self.hyper_analysis(y_abs)
        return y, z, z2_hw

    def hyper_synthesis(self, z_hat_bf16, mid_hw, y_hw):
        """h_s: z_hat -> (scales, None) (fp32, ReLU'd as models/google.py:275-282); mean-scale: -> (scales, means), the two
        channel halves of the last layer's output (`chunk(2, 1)`, models/google.py:397-398)."""
        t = self.h_s["0"](z_hat_bf16, relu=self._act, out_hw=mid_hw)
        t = self.h_s["2"](t, relu=self._act, out_hw=y_hw)
        if not self.spec.mean_scale:
            scales, _ = self.h_s["4"](t, relu=True, latent_head=True)
            return scales, None
        params, _ = self.h_s["4"](t, latent_head="plain")
        scales, means = params.chunk(2, 1)            # channel halves of an NHWC tensor: re-homed for the flat kernels
        return scales.contiguous(memory_format=Fn.CL), means.contiguous(memory_format=Fn.CL)

    def synthesis(self, y_hat_bf16, level_hw):
        """g_s: y_hat -> the reconstructed levels (level 0 first), each cropped to its level's H x W."""
        s, tags = self.g_s, self._tags
        n = len(tags)
        self._gate(f"g_s.up{tags[-1]}.weight")
        hs = Fn.SideStreams.heads if (self.training and torch.is_grad_enabled()) else None
        main = torch.cuda.current_stream()

        def produced():
            """Event right behind the kernel that produced d_l: the head depends on d_l only, not on whatever the main
            stream records between d_l and the head's own (later) recording."""
            if hs is None:
                return None
            ev = torch.cuda.Event()
            ev.record(main)
            return ev

        def head(layer, d, link, ev):
            if hs is None:
                return layer(d, link=link, side=True)
            d.record_stream(hs)
            with torch.cuda.stream(hs):
                hs.wait_event(ev)
                out = layer(d, link=link, side=True)
            out.record_stream(main)
            return out

        # Each d_l feeds the next up-sampling stage (main consumer) and its 3x3 head (side consumer).  The main consumer is
        # recorded FIRST: autograd then runs the head's backward first, which parks its data gradient in the GDNLink, and
        # the up-sampling stage's data-gradient kernel adds it and runs the IGDN's backward in its epilogue.
        outs = [None] * n
        d, link, pending = y_hat_bf16, None, None
        for i in range(n - 1, 0, -1):
            t = tags[i]
            d, lk = getattr(s, f"up{t}").forward_gdn(getattr(s, f"igdn{t}"), d, out_hw=level_hw[i], link=link)
            if lk is not None:
                lk.expect_side = True
            ev = produced()
            if pending is not None:               # the previous level's head, recorded after this level's main consumer
                outs[pending[0]] = head(*pending[1:])
            pending = (i, getattr(s, f"head{t}"), d, lk, ev)
            link = lk
        outs[0] = getattr(s, f"up{tags[0]}")(d, out_hw=level_hw[0], link=link)
        outs[pending[0]] = head(*pending[1:])
        if hs is not None:
            main.wait_stream(hs)
        return tuple(outs)

    # ------------------------------------------------------------------ entropy-coded path (models/google.py:313-333)
    @torch.no_grad()
    def compress(self, feats):
        """ScaleHyperprior.compress: -> {"strings": [y_strings, z_strings], "shape": z H x W, plus the sizes the
        ragged pyramid needs to be rebuilt ("y_hw", "mid_hw", "level_hw")}.  The scales that index y's CDFs are
        computed from the DECODED z_hat, by the same kernels the decoder runs, so both sides see identical bits."""
        if self.entropy_bottleneck._offset.numel() == 0 or self.gaussian_conditional._offset.numel() == 0:
            raise ValueError("Uninitialized CDFs. Run update() first")
        lv = self.spec.levels
        ps = tuple(Fn.nhwc(feats[k], Fn.BF16) for k in lv)
        y, z, mid_hw = self.analysis(ps)
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        if getattr(self, "coder_lanes_auto", False):     # ~1024 symbols per string: <= ~4 % container overhead
            eb.coder_lanes = max(1, min(256, z[0].numel() // 1024))
            gc.coder_lanes = max(1, min(256, y[0].numel() // 1024))
        z_strings = eb.compress(z)
        z_hat = eb.decompress(z_strings, z.shape[-2:])
        with L.batch_invariant():     # same bits whatever the batch: the decoder may see the frames one at a time
            scales, means = self.hyper_synthesis(Fn.nhwc(z_hat.contiguous(), Fn.BF16), mid_hw, y.shape[2:])
        indexes = gc.build_indexes(scales)
        y_strings = gc.compress(y, indexes, means=means)
        return {"strings": [y_strings, z_strings], "shape": tuple(z.shape[-2:]), "y_hw": tuple(y.shape[2:]),
                "mid_hw": tuple(mid_hw), "level_hw": [tuple(p.shape[2:]) for p in ps]}

    @torch.no_grad()
    def decompress(self, strings, shape, y_hw, mid_hw, level_hw):
        """ScaleHyperprior.decompress: -> {"x_hat": {level: tensor}}."""
        assert isinstance(strings, list) and len(strings) == 2
        eb, gc = self.entropy_bottleneck, self.gaussian_conditional
        if getattr(self, "coder_lanes_auto", False):     # lane counts are read from the strings
            eb.coder_lanes = gc.coder_lanes = 0
        z_hat = eb.decompress(strings[1], shape)
        with L.batch_invariant():
            scales, means = self.hyper_synthesis(Fn.nhwc(z_hat.contiguous(), Fn.BF16), tuple(mid_hw), tuple(y_hw))
        indexes = gc.build_indexes(scales)
        y_hat = gc.decompress(strings[0], indexes, z_hat.dtype, means=means)
        x_hat = self.synthesis(Fn.nhwc(y_hat.contiguous(), Fn.BF16), [tuple(h) for h in level_hw])
        return {"x_hat": dict(zip(self.spec.levels, x_hat))}

    # ------------------------------------------------------------------ CompressionModel contract
    def aux_loss(self):
        """models/base.py:115-144: sum of the EntropyBottleneck aux losses."""
        return sum(m.loss() for m in self.modules() if isinstance(m, EntropyBottleneck))

    def update(self, scale_table=None, force=False):
        """models/base.py:90-113."""
        if scale_table is None:
            scale_table = get_scale_table()
        updated = self.gaussian_conditional.update_scale_table(scale_table, force=force)
        for m in self.modules():
            if isinstance(m, EntropyBottleneck):
                updated |= m.update(force=force)
        return updated

    def load_state_dict(self, state_dict, strict=True):
        """models/base.py:67-88: CDF buffers are empty until update(); resize them to the checkpoint."""
        for name, module in self.named_modules():
            if not isinstance(module, (EntropyBottleneck, GaussianConditional)):
                continue
            buffers = ["_quantized_cdf", "_offset", "_cdf_length"]
            if isinstance(module, GaussianConditional):
                buffers.append("scale_table")
            for b in buffers:
                key = f"{name}.{b}"
                if key in state_dict:
                    src = state_dict[key]
                    cur = getattr(module, b)
                    if cur.shape != src.shape:
                        setattr(module, b, torch.empty(src.shape, dtype=cur.dtype, device=cur.device))
        out = super().load_state_dict(state_dict, strict=strict)
        self.refresh_shadows()
        return out

    # ------------------------------------------------------------------ arena / shadows
    def conv_layers(self):
        return [m for m in self.modules() if isinstance(m, _ConvBase)]

    def build_arena(self, alloc=None):
        """Move all parameters (and .grad) into flat fp32 arenas.  `alloc(numel, dtype, device)` (optional) provides the
        three big allocations -- parameters, gradients, bf16 operands -- e.g. from peer-mapped memory for the
        data-parallel exchange (parallel.symmetric_alloc); calling it again re-homes the parameters."""
        self.arena = ParamArena(self, alloc=alloc)
        return self.arena

    def refresh_shadows(self, cast=True):
        """Rebuild the bf16 GEMM operands after the fp32 weights changed (optimizer step / load)."""
        if self.arena is not None:
            self.arena.refresh_shadows(cast=cast)
        else:
            for m in self.conv_layers():
                if m.weight.is_cuda:
                    m._state.refresh(m.weight, force=True)


def _round_up(n, k):
    return (n + k - 1) // k * k


class ParamArena:
    """Flat fp32 storage for parameters / gradients (+ one bf16 shadow block).

    "net" holds every parameter except `*.quantiles`, "aux" holds those -- the split of
    compressai/optimizers/net_aux.py:46-57.  Every parameter keeps its logical shape (conv weights
    stay GEMM-layout permuted views), only its storage moves.  Physical order of the net arena:

        [GDN beta/gamma of every layer] [EntropyBottleneck blocks, C x 58 interleaved] [the rest]

    * the EntropyBottleneck's 14 reference-named parameters become strided views of the per-channel
      block of 58 floats the kernels read, and their .grad of the block the backward kernel writes;
    * the GDN group is reparametrised (forward) and its gradient mapped back (backward) by ONE
      batched launch each per step; its "effective" gradient scratch sits behind the parameter
      gradients in the same allocation so one fill zeroes everything.
    """

    ALIGN = 64   # floats: 256-byte aligned slices (TMA bases, float4 kernels)

    def __init__(self, model, alloc=None):
        if alloc is None:
            alloc = lambda numel, dtype, device: torch.zeros(numel, dtype=dtype, device=device)
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        if not named or not named[0][1].is_cuda:
            raise RuntimeError("build_arena(): move the model to a CUDA device first (no CPU fallback)")
        dev = named[0][1].device
        self.device = dev
        self.model = model
        name_of = {id(p): n for n, p in named}
        gdns = [m for m in model.modules() if isinstance(m, GDN)]
        ebs = [m for m in model.modules() if isinstance(m, EntropyBottleneck)]
        gdn_ids = {id(p) for m in gdns for p in (m.beta, m.gamma)}
        eb_ids = {id(getattr(m, pn)) for m in ebs for pn, _, _ in m.PACK_COLUMNS}
        rest = [(n, p) for n, p in named if id(p) not in gdn_ids and id(p) not in eb_ids and not n.endswith(".quantiles")]
        # the parameters kernels read in fp32 (biases ...) in front of the conv weights (read as bf16 operands only): the
        # data-parallel exchange publishes [0, first weight) in fp32 and the rest as bf16 (csrc/exchange.cu)
        rest = [(n, p) for n, p in rest if p.dim() != 4] + [(n, p) for n, p in rest if p.dim() == 4]
        self.aux = [(n, p) for n, p in named if n.endswith(".quantiles")]
        self.offsets = {}
        # ---- plan the net arena
        off = 0
        plan = []            # (name, param, offset, shape, stride)
        gdn_plan = []
        for m in gdns:
            for p in (m.beta, m.gamma):
                plan.append((name_of[id(p)], p, off, p.shape, p.stride()))
                gdn_plan.append((m, p, off))
                off += _round_up(p.numel(), self.ALIGN)
        eb_blocks = []
        for m in ebs:
            C, W = m.channels, m.PACK_WIDTH
            eb_blocks.append((m, off))
            for pn, a, b in m.PACK_COLUMNS:
                p = getattr(m, pn)
                cols = p.shape[2]
                plan.append((name_of[id(p)], p, off + a, p.shape, (W, cols, 1)))
            off += _round_up(C * W, self.ALIGN)
        for n, p in rest:
            plan.append((n, p, off, p.shape, p.stride()))
            off += _round_up(p.numel(), self.ALIGN)
        n_net = max(off, self.ALIGN)
        # ---- effective-gradient scratch of the GDN group behind the parameter gradients
        g_off, o32, o16 = n_net, 0, 0
        rows, bounds = [], []
        for m, p, poff in gdn_plan:
            is_beta = p is m.beta
            rows.append([poff, p.numel(), o32 if is_beta else -1, -1 if is_beta else o16, g_off - 0])
            bounds.append(m._state.beta_bound if is_beta else m._state.gamma_bound)
            if is_beta:
                o32 += _round_up(p.numel(), self.ALIGN)
            else:
                o16 += _round_up(p.numel(), self.ALIGN)
            g_off += _round_up(p.numel(), self.ALIGN)
        with torch.no_grad():
            self.net_params = alloc(n_net, torch.float32, dev)
            self._grads_all = alloc(g_off, torch.float32, dev)
            self.net_grads = self._grads_all[:n_net]
            self.net = []
            for n, p, o, shape, stride in plan:
                self.offsets[n] = o
                view = self.net_params.as_strided(shape, stride, o)
                view.copy_(p.data)
                p.data = view
                p.grad = self.net_grads.as_strided(shape, stride, o)
                self.net.append((n, p))
            aoff = 0
            for n, p in self.aux:
                self.offsets[n] = aoff
                aoff += _round_up(p.numel(), self.ALIGN)
            self.aux_params = torch.zeros(max(aoff, self.ALIGN), dtype=torch.float32, device=dev)
            self.aux_grads = torch.zeros_like(self.aux_params)
            for n, p in self.aux:
                o = self.offsets[n]
                view = self.aux_params.as_strided(p.shape, p.stride(), o)
                view.copy_(p.data)
                p.data = view
                p.grad = self.aux_grads.as_strided(p.shape, p.stride(), o)
        for m, o in eb_blocks:
            C, W = m.channels, m.PACK_WIDTH
            m._packed = self.net_params[o:o + C * W].view(C, W)
            m._packed_grad = self.net_grads[o:o + C * W].view(C, W)
        # ---- GDN group: effective values, gradient scratch, segment table
        self.gdn_layers = gdns
        self.eff32 = torch.zeros(max(o32, 1), dtype=torch.float32, device=dev)
        self.eff16 = torch.zeros(max(o16, 1), dtype=torch.bfloat16, device=dev)
        self.gdn_table = torch.tensor(rows, dtype=torch.int64, device=dev).reshape(-1, 5)
        self.gdn_bounds = torch.tensor(bounds, dtype=torch.float32, device=dev)
        self.gdn_max_n = max([r[1] for r in rows], default=0)
        for (m, p, poff), r in zip(gdn_plan, rows):
            st, C = m._state, m.in_channels
            st.arena = self
            if p is m.beta:
                st.beta_eff = self.eff32[r[2]:r[2] + C]
                st.g_beta_eff = self._grads_all[r[4]:r[4] + C]
            else:
                st.gamma_bf = self.eff16[r[3]:r[3] + C * C].view(C, C)
                st.g_gamma_eff = self._grads_all[r[4]:r[4] + C * C].view(C, C)
        self._gdn_versions = None
        self._finish_queued = False
        self.gates = {}            # parameter name -> event of the optimiser slice starting there (train.TrainStep)
        self.n_net = n_net
        # ---- bf16 GEMM operands of the conv weights: slices of one shadow of the whole arena
        self.net_bf16 = alloc(self.net_params.numel(), torch.bfloat16, dev)
        for m in model.conv_layers():
            o = self.offsets[name_of[id(m.weight)]]
            st = m._state
            st.w_direct = self.net_bf16[o:o + m.weight.numel()].view(st.D0, st.T, st.D1)
        self.refresh_shadows(cast=True)

    def names(self, group="net"):
        return [n for n, _ in (self.net if group == "net" else self.aux)]

    # ------------------------------------------------------------------ derived operands
    def refresh_shadows(self, cast=True):
        """bf16 GEMM operands of all conv weights = one cast of the flat arena, and the GDN group's
        effective beta / gamma (Adam writes the bf16 block in its own pass and calls prepare())."""
        if cast:
            L.cast_f32_bf16(ptr(self.net_params), ptr(self.net_bf16), self.net_params.numel(), stream())
        for m in self.model.conv_layers():
            m._state._version = m.weight._version
        self.prepare()

    def _versions(self):
        return tuple(p._version for m in self.gdn_layers for p in (m.beta, m.gamma))

    def prepare(self):
        """NonNegativeParametrizer forward of every GDN layer, one launch (ops/parametrizers.py:60-63)."""
        if self.gdn_layers:
            L.reparam_forward_multi(ptr(self.net_params), ptr(self.gdn_table), ptr(self.gdn_bounds), self.gdn_table.shape[0],
                                    self.gdn_max_n, Fn.PEDESTAL, ptr(self.eff32), ptr(self.eff16), stream())
        self._gdn_versions = self._versions()

    def ensure_prepared(self, beta, gamma):
        if self._gdn_versions != self._versions():      # someone wrote the parameters through torch
            self.prepare()

    def queue_finish_backward(self):
        if not self._finish_queued:
            self._finish_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self.finish_backward)

    def finish_backward(self):
        """End of a backward pass: map the GDN group's effective gradients onto beta / gamma
        (LowerBound gradient rule, ops/bound_ops.py:40-42), one launch; then report them ready."""
        self._finish_queued = False
        if not self.gdn_layers:
            return
        Fn.SideStreams.join_wgrad()              # gamma / beta gradient GEMMs ran on the weight-gradient streams
        L.reparam_backward_multi(ptr(self._grads_all), ptr(self.net_params), ptr(self.gdn_table), ptr(self.gdn_bounds),
                                 self.gdn_table.shape[0], self.gdn_max_n, ptr(self.net_grads), stream())
        for m in self.gdn_layers:
            if m._state.grad_ready is not None:
                m._state.grad_ready([m.beta, m.gamma])

    def zero_grad(self):
        self._grads_all.zero_()
        self.aux_grads.zero_()
