"""The frozen L-MSFC topology (ONE spec file).

DESIGN DECISION, NOT A REFERENCE FACT: the reference checkout ships no L-MSFC
model.  north_star fixes the ingredients (P2-P5 fusion encoder
and decoder from strided conv + GDN/IGDN, scale hyperprior, EntropyBottleneck,
GaussianConditional, bpp + lambda*MSE); this file fixes how they are wired.
Both the product model (model.py) and the oracle (oracle/lmsfc.py) are built
from it, and so is the FLOP/byte counter used by bench.py.

    enc:  e2 = GDN(conv5s2(F -> N)(p2))                    @ /8
          e3 = GDN(conv5s2(N+F -> N)(cat[e2, p3]))         @ /16
          e4 = GDN(conv5s2(N+F -> N)(cat[e3, p4]))         @ /32
          y  =     conv5s2(N+F -> M)(cat[e4, p5])          @ /64
    hyp:  z = h_a(|y|) ; z_hat, Lz = EntropyBottleneck(N)(z) ; s = h_s(z_hat)
          y_hat, Ly = GaussianConditional(None)(y, s)      (models/google.py:267-303)
          hyperprior="mean-scale": z = h_a(y) ; (s, mu) = chunk(h_s(z_hat)) ; y_hat, Ly = GC(y, s, means=mu)  (:337-405)
    dec:  d5 = IGDN(deconv5s2(M -> N)(y_hat)) @ /32 ; p5_hat = conv3(N -> F)(d5)
          d4 = IGDN(deconv5s2(N -> N)(d5))    @ /16 ; p4_hat = conv3(N -> F)(d4)
          d3 = IGDN(deconv5s2(N -> N)(d4))    @ /8  ; p3_hat = conv3(N -> F)(d3)
          p2_hat =  deconv5s2(N -> F)(d3)     @ /4
    loss: bpp over {y, z}, num_pixels = B * (4*H2) * (4*W2) ;
          D = sum_l w_l * mean((p_l_hat - p_l)^2) ; loss = bpp + lambda*255^2*D
          (losses/rate_distortion.py:47-59 semantics kept).
    RNG order in one forward: EB noise ([C,1,B*h*w] layout) then GC noise
    (models/google.py:295-297).

Odd sizes: every deconv output is cropped to the target level's H x W.
"""
from dataclasses import dataclass
from typing import Tuple


@dataclass(frozen=True)
class LMSFCSpec:
    """One frozen topology.  The default is the 4-level, 256-channel Faster / Mask R-CNN FPN pyramid (P2..P5); any number
    of levels >= 2 with per-level channel counts builds the same pattern (JDE_SPEC: the 3-scale, 256/512/1024-channel
    split of the JDE tracker, cfgs/vision_model/default.yaml:33, compressai_vision/model_wrappers/jde.py:62-180):

        enc:  e_0 = GDN(conv5s2(F_0 -> N)(x_0)) ; e_l = GDN(conv5s2(N+F_l -> N)(cat[e_{l-1}, x_l])) ;
              y = conv5s2(N+F_last -> M)(cat[e_{last-1}, x_last])
        dec:  d_l = IGDN(deconv5s2(. -> N)(.)) for l = last .. 1 with head conv3(N -> F_l) ; x_0_hat = deconv5s2(N -> F_0)(d_1)
    """
    F: int = 256            # feature channels of level 0 (all levels unless level_channels says otherwise)
    N: int = 256            # hidden / hyper-latent channels
    M: int = 256            # latent channels
    levels: Tuple[str, ...] = ("p2", "p3", "p4", "p5")
    level_weights: Tuple[float, ...] = (1.0, 1.0, 1.0, 1.0)
    distortion_scale: float = 255.0 ** 2
    scale_bound: float = 0.11
    likelihood_bound: float = 1e-9
    level_channels: Tuple[int, ...] = ()     # () -> F for every level
    base_stride: int = 4                     # input pixels per level-0 feature pixel (P2: 4; JDE's first split: 8)
    # "scale": models/google.py:267-303 (h_a(|y|), ReLU, h_s -> scales).  "mean-scale": models/google.py:337-405 -- h_a(y),
    # LeakyReLU, h_s: deconv(N -> M), deconv(M -> hyper_mid), conv3(hyper_mid -> 2M), chunk -> (scales, means).  The
    # reference's middle width 3M/2 (= 384) is not a multiple of the 256-channel weight-gradient tile: default 2M.
    hyperprior: str = "scale"
    hyper_mid: int = 0                       # 0 -> 2 * M

    def __post_init__(self):
        if len(self.levels) < 2 or len(self.level_weights) != len(self.levels):
            raise ValueError("LMSFCSpec: at least two levels, one weight per level")
        if self.level_channels and len(self.level_channels) != len(self.levels):
            raise ValueError("LMSFCSpec: one channel count per level")
        if self.hyperprior not in ("scale", "mean-scale"):
            raise ValueError("LMSFCSpec: hyperprior is 'scale' or 'mean-scale'")

    @property
    def mean_scale(self):
        return self.hyperprior == "mean-scale"

    @property
    def mid(self):
        return self.hyper_mid or 2 * self.M

    @property
    def channels(self):
        return tuple(self.level_channels) if self.level_channels else (self.F,) * len(self.levels)

    def tag(self, i):
        """Module-name suffix of level i: the digits of its key ("p3" -> "3", "l61" -> "61")."""
        d = "".join(ch for ch in self.levels[i] if ch.isdigit())
        return d or str(i)

    def gate_names(self):
        """First parameter of every optimiser slice of the pipelined train step, in arena (= forward) order."""
        n = len(self.levels)
        return tuple(f"g_a.conv{self.tag(i)}.weight" for i in range(1, n)) + ("h_a.0.weight", f"g_s.up{self.tag(n - 1)}.weight")

    def level_shapes(self, H2: int, W2: int):
        """(H, W) of every level given level 0's size: each level halves, rounding up
        (detectron2 pads inputs to a multiple of 32, so bench shapes are exact)."""
        out, h, w = [], H2, W2
        for _ in self.levels:
            out.append((h, w))
            h, w = (h + 1) // 2, (w + 1) // 2
        return out

    def num_pixels(self, B: int, H2: int, W2: int) -> int:
        return B * (self.base_stride * H2) * (self.base_stride * W2)

    def conv_layers(self, H2: int, W2: int):
        """[(name, kind, k, Cin, Cout, Hin, Win, Hout, Wout)] for the FLOP counter.
        kind: 'conv' (stride 2 for k=5, 1 for k=3), 'deconv' (k5 s2), 'gdn'."""
        N, M, ch = self.N, self.M, self.channels
        shp = self.level_shapes(H2, W2)
        c = lambda v: (v + 1) // 2
        n = len(self.levels)
        L = []
        for i in range(n):
            h, w = shp[i]
            cin = ch[i] + (N if i else 0)
            last = i == n - 1
            L.append((f"g_a.conv{self.tag(i)}", "conv", 5, cin, M if last else N, h, w, c(h), c(w)))
            if not last:
                L.append((f"g_a.gdn{self.tag(i)}", "gdn", 1, N, N, c(h), c(w), c(h), c(w)))
        yy = (c(shp[-1][0]), c(shp[-1][1]))
        z1 = (c(yy[0]), c(yy[1])); z2 = (c(z1[0]), c(z1[1]))
        L += [("h_a.0", "conv", 3, M, N, *yy, *yy), ("h_a.2", "conv", 5, N, N, *yy, *z1), ("h_a.4", "conv", 5, N, N, *z1, *z2)]
        if self.mean_scale:
            L += [("h_s.0", "deconv", 5, N, M, *z2, *z1), ("h_s.2", "deconv", 5, M, self.mid, *z1, *yy),
                  ("h_s.4", "conv", 3, self.mid, 2 * M, *yy, *yy)]
        else:
            L += [("h_s.0", "deconv", 5, N, N, *z2, *z1), ("h_s.2", "deconv", 5, N, N, *z1, *yy), ("h_s.4", "conv", 3, N, M, *yy, *yy)]
        src = yy
        for i in range(n - 1, 0, -1):
            h, w = shp[i]
            L.append((f"g_s.up{self.tag(i)}", "deconv", 5, M if i == n - 1 else N, N, *src, h, w))
            L.append((f"g_s.igdn{self.tag(i)}", "gdn", 1, N, N, h, w, h, w))
            L.append((f"g_s.head{self.tag(i)}", "conv", 3, N, ch[i], h, w, h, w))
            src = (h, w)
        L.append((f"g_s.up{self.tag(0)}", "deconv", 5, N, ch[0], *src, *shp[0]))
        return L

    def flops_fwd(self, H2: int, W2: int) -> int:
        """sum 2*k^2*Cin*Cout*(out px for conv | in px for deconv)
        + sum_GDN 2*C^2*H*W, per feature-set."""
        total = 0
        for (_, kind, k, ci, co, hi, wi, ho, wo) in self.conv_layers(H2, W2):
            if kind == "deconv":
                total += 2 * k * k * ci * co * hi * wi
            else:
                total += 2 * k * k * ci * co * ho * wo
        return total

    def flops_train(self, H2: int, W2: int) -> int:
        return 3 * self.flops_fwd(H2, W2)


DEFAULT_SPEC = LMSFCSpec()
# BASELINE configs[4]: JDE-1088x608, the three Darknet-53 split tensors (layers 36 / 61 / 74: 256 x 76 x 136,
# 512 x 38 x 68, 1024 x 19 x 34; cfgs/vision_model/default.yaml:33, data/mpeg-fcm/MPEGTVDTRACKING/fpn-sizes)
JDE_SPEC = LMSFCSpec(levels=("l36", "l61", "l74"), level_weights=(1.0, 1.0, 1.0), level_channels=(256, 512, 1024), base_stride=8)

# the same pyramid under the mean-scale hyperprior of models/google.py:337-405
MEAN_SCALE_SPEC = LMSFCSpec(hyperprior="mean-scale")

# first six reference MSE lambdas, compressai/docs/source/zoo.rst:68
LAMBDAS = (0.0018, 0.0035, 0.0067, 0.0130, 0.0250, 0.0483)
