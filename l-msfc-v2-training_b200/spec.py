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
from dataclasses import dataclass, field
from typing import Tuple


@dataclass(frozen=True)
class LMSFCSpec:
    F: int = 256            # FPN feature channels (P2..P5)
    N: int = 256            # hidden / hyper-latent channels
    M: int = 256            # latent channels
    levels: Tuple[str, ...] = ("p2", "p3", "p4", "p5")
    level_weights: Tuple[float, ...] = (1.0, 1.0, 1.0, 1.0)
    distortion_scale: float = 255.0 ** 2
    scale_bound: float = 0.11
    likelihood_bound: float = 1e-9

    def level_shapes(self, H2: int, W2: int):
        """(H, W) of P2..P5 given P2's size: each level halves, rounding up
        (detectron2 pads inputs to a multiple of 32, so bench shapes are exact)."""
        out, h, w = [], H2, W2
        for _ in self.levels:
            out.append((h, w))
            h, w = (h + 1) // 2, (w + 1) // 2
        return out

    def num_pixels(self, B: int, H2: int, W2: int) -> int:
        return B * (4 * H2) * (4 * W2)

    def conv_layers(self, H2: int, W2: int):
        """[(name, kind, k, Cin, Cout, Hin, Win, Hout, Wout)] for the FLOP counter.
        kind: 'conv' (stride 2 for k=5, 1 for k=3), 'deconv' (k5 s2), 'gdn'."""
        F, N, M = self.F, self.N, self.M
        (h2, w2), (h3, w3), (h4, w4), (h5, w5) = self.level_shapes(H2, W2)
        c = lambda v: (v + 1) // 2
        e2 = (c(h2), c(w2)); e3 = (c(h3), c(w3)); e4 = (c(h4), c(w4)); yy = (c(h5), c(w5))
        z1 = (c(yy[0]), c(yy[1])); z2 = (c(z1[0]), c(z1[1]))
        L = [
            ("g_a.conv2", "conv", 5, F, N, h2, w2, *e2), ("g_a.gdn2", "gdn", 1, N, N, *e2, *e2),
            ("g_a.conv3", "conv", 5, N + F, N, h3, w3, *e3), ("g_a.gdn3", "gdn", 1, N, N, *e3, *e3),
            ("g_a.conv4", "conv", 5, N + F, N, h4, w4, *e4), ("g_a.gdn4", "gdn", 1, N, N, *e4, *e4),
            ("g_a.conv5", "conv", 5, N + F, M, h5, w5, *yy),
            ("h_a.0", "conv", 3, M, N, *yy, *yy),
            ("h_a.2", "conv", 5, N, N, *yy, *z1),
            ("h_a.4", "conv", 5, N, N, *z1, *z2),
            ("h_s.0", "deconv", 5, N, N, *z2, *z1),
            ("h_s.2", "deconv", 5, N, N, *z1, *yy),
            ("h_s.4", "conv", 3, N, M, *yy, *yy),
            ("g_s.up5", "deconv", 5, M, N, *yy, h5, w5), ("g_s.igdn5", "gdn", 1, N, N, h5, w5, h5, w5),
            ("g_s.head5", "conv", 3, N, F, h5, w5, h5, w5),
            ("g_s.up4", "deconv", 5, N, N, h5, w5, h4, w4), ("g_s.igdn4", "gdn", 1, N, N, h4, w4, h4, w4),
            ("g_s.head4", "conv", 3, N, F, h4, w4, h4, w4),
            ("g_s.up3", "deconv", 5, N, N, h4, w4, h3, w3), ("g_s.igdn3", "gdn", 1, N, N, h3, w3, h3, w3),
            ("g_s.head3", "conv", 3, N, F, h3, w3, h3, w3),
            ("g_s.up2", "deconv", 5, N, F, h3, w3, h2, w2),
        ]
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

# first six reference MSE lambdas, compressai/docs/source/zoo.rst:68
LAMBDAS = (0.0018, 0.0035, 0.0067, 0.0130, 0.0250, 0.0483)
