"""RateDistortionLoss for the multi-scale codec (compressai/losses/rate_distortion.py:38-59).

loss = lmbda * 255^2 * D + bpp with bpp = sum_k sum(ln L_k) / (-ln 2 * num_pixels) and
D = sum_l w_l * mean((x_hat_l - x_l)^2) (level weights and num_pixels from spec.py).  The MSE value
and its gradient come out of one fused pass per level; sum(ln L) was already reduced inside the
likelihood kernels.
"""
import torch.nn as nn

from . import functional as Fn
from .spec import DEFAULT_SPEC


class RateDistortionLoss(nn.Module):
    def __init__(self, lmbda=0.01, metric="mse", return_type="all", spec=DEFAULT_SPEC):
        super().__init__()
        if metric != "mse":
            raise NotImplementedError(f"{metric} is not implemented!")   # rate_distortion.py:41-46 (ms-ssim is image-only)
        self.lmbda = float(lmbda)
        self.return_type = return_type
        self.spec = spec

    def forward(self, output, target, extra=None):
        """extra (optional device scalar, e.g. the aux loss): rides along in out["scalars"] = {loss, bpp, mse, extra},
        written by the loss kernel itself -- a train step reads its four numbers back as one 16-byte vector."""
        lv = self.spec.levels
        x_hat = output["x_hat"]
        nhwc_in = output.get("inputs_nhwc")
        targets = [nhwc_in[k] if nhwc_in is not None else target[k] for k in lv]
        B, _, H2, W2 = targets[0].shape
        num_pixels = self.spec.num_pixels(B, H2, W2)
        loss, bpp, mse, s4 = Fn.rd_loss_apply([x_hat[k] for k in lv], targets, list(output["likelihoods"].values()),
                                              self.lmbda * self.spec.distortion_scale, num_pixels, self.spec.level_weights,
                                              extra=extra, with_scalars=True)
        out = {"bpp_loss": bpp, "mse_loss": mse, "loss": loss}
        if s4 is not None and extra is not None:
            out["scalars"] = s4
        return out if self.return_type == "all" else out[self.return_type]
