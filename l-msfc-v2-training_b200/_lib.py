"""ctypes binding of libb200lmsfc.so (C ABI in include/b200lmsfc.h).

There is no CPU fallback: importing this module without the built library raises, and every call
that fails raises with the library's message (ValueError for the reference's ValueError cases,
RuntimeError for CUDA errors).
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int64, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# LMSFC_LIB selects an alternative build of the same library (e.g. the TIMELINE=1 debug build)
LIB_PATH = os.environ.get("LMSFC_LIB") or os.path.join(_HERE, "libb200lmsfc.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C l-msfc-v2-training_b200/csrc`).  This package has no CPU / PyTorch fallback.")

lib = ctypes.CDLL(LIB_PATH)


class ConvDesc(Structure):
    _fields_ = [("kind", c_int), ("k", c_int), ("stride", c_int), ("B", c_int), ("Hin", c_int), ("Win", c_int),
                ("Hout", c_int), ("Wout", c_int), ("c0", c_int), ("c1", c_int), ("Cout", c_int)]


MAX_RANKS, XCHG_SLOTS = 8, 32


class XchgPeers(Structure):
    """lmsfc_xchg_peers: the same buffer of every rank, peer-mapped (include/b200lmsfc.h)."""
    _fields_ = [("world", c_int), ("rank", c_int), ("grads", c_void_p * MAX_RANKS), ("params", c_void_p * MAX_RANKS),
                ("shadow", c_void_p * MAX_RANKS), ("flags", c_void_p * MAX_RANKS), ("norm_all", c_void_p * MAX_RANKS),
                ("tick", c_void_p), ("mc_grads", c_void_p), ("mc_params", c_void_p), ("mc_shadow", c_void_p)]


P = c_void_p
_SIGS = {
    "lmsfc_conv_forward": [POINTER(ConvDesc), P, P, P, P, c_int, P, P, P, P],
    "lmsfc_conv_dgrad": [POINTER(ConvDesc), P, P, c_int, P, P],
    "lmsfc_conv_wgrad": [POINTER(ConvDesc), P, P, P, P, P, P],
    "lmsfc_conv_gdn_fusable": [POINTER(ConvDesc), c_int, c_int, POINTER(c_int)],
    "lmsfc_conv_gdn_forward": [POINTER(ConvDesc), P, P, P, P, P, P, c_int, P, P, P, P, P],
    "lmsfc_conv_dgrad_gdn_backward": [POINTER(ConvDesc), P, P, c_int, P, P, P, P, c_int, P, P, P],
    "lmsfc_gdn_forward": [P, P, P, c_int, c_int64, c_int, P, P, P, P],
    "lmsfc_gdn_backward": [P, P, P, P, c_int, c_int64, c_int, P, c_int, P, P, P, P, P],
    "lmsfc_gdn_wgrad": [P, P, c_int64, c_int, P, P, P],
    "lmsfc_reparam_forward": [P, c_int64, c_float, c_float, P, P, P],
    "lmsfc_reparam_backward": [P, P, c_int64, c_float, P, P],
    "lmsfc_reparam_forward_multi": [P, P, P, c_int, c_int64, c_float, P, P, P],
    "lmsfc_reparam_backward_multi": [P, P, P, P, c_int, c_int64, P, P],
    "lmsfc_gc_forward": [P, P, P, P, c_int64, c_int, c_float, c_float, c_uint64, c_uint64, P, P, P, P, P, P],
    "lmsfc_gc_backward": [P, P, P, P, P, P, c_int64, c_int, c_float, c_float, c_float, P, P, P, P],
    "lmsfc_eb_forward": [P, P, P, P, c_int64, c_int, c_int, c_float, c_uint64, c_uint64, P, P, P, P, P, P],
    "lmsfc_eb_backward": [P, P, P, P, P, c_int64, c_int, c_int, c_float, c_float, P, P, P],
    "lmsfc_eb_aux_loss": [P, P, P, c_int, P, P, P],
    "lmsfc_rng_advance": [P, P],
    "lmsfc_eb_pmf": [P, P, c_int, c_int, P, c_int, P, P],
    "lmsfc_gc_pmf": [P, P, c_int, c_int, P, c_int, P, P],
    "lmsfc_mse_forward_backward": [P, P, c_int64, c_float, P, P, P],
    "lmsfc_mse_forward_backward_multi": [c_int, P, P, P, P, P, P, P],
    "lmsfc_rd_loss": [c_int, P, P, P, P, P, c_int, P, c_float, P, c_float, P, P, P],
    "lmsfc_loss_finalize": [P, P, c_int, c_int, c_float, P, P],
    "lmsfc_relu_backward": [P, P, c_int64, P],
    "lmsfc_leaky_relu_backward": [P, P, c_int64, c_float, P],
    "lmsfc_abs_backward": [P, P, P, c_int64, c_int, P, P],
    "lmsfc_colsum": [P, c_int64, c_int, P, P],
    "lmsfc_colsum_static": [P, c_int64, c_int, P, P],
    "lmsfc_cast_f32_bf16": [P, P, c_int64, P],
    "lmsfc_cast_bf16_f32": [P, P, c_int64, P],
    "lmsfc_nchw_to_nhwc_bf16": [P, P, c_int, c_int, c_int, P],
    "lmsfc_sumsq": [P, c_int64, P, P, P],
    "lmsfc_sumsq_part": [P, c_int64, P, P],
    "lmsfc_sum_parts": [P, c_int, P, P, P],
    "lmsfc_sumsq_part_final": [P, c_int64, P, c_int, c_int, P, P, P],
    "lmsfc_adam_step": [P, P, P, P, P, c_int64, c_float, c_float, c_float, c_float, c_float, P, P, P, P],
    "lmsfc_adam_prep": [c_float, P, c_float, c_float, c_float, P, P, P, P, c_int, P, P],
    "lmsfc_adam_update": [P, P, P, P, P, c_int64, c_float, c_float, c_float, P, c_int, P],
    "lmsfc_pmf_to_quantized_cdf": [P, c_int, c_int, P],
    "lmsfc_pmf_rows_to_cdf": [P, c_int, P, P, c_int, c_int, P, c_int, P, P],
    "lmsfc_rans_encode": [P, P, c_int, c_int64, c_int, P, c_int, c_int, P, P, P, c_int64, P, P, P, c_int64, P, P],
    "lmsfc_rans_decode": [P, c_int64, P, P, c_int, c_int64, c_int, P, c_int, c_int, P, P, P, P, P],
    "lmsfc_rans_encode_indexed": [P, P, c_int, c_int64, c_int, P, c_int, c_int, P, P, P, c_int64, P, P, P, c_int64, P,
                                  c_int, c_int, P, P],
    "lmsfc_rans_decode_indexed": [P, c_int64, P, P, c_int, c_int64, c_int, P, c_int, c_int, P, P, P, P, c_int, c_int, P, P],
    "lmsfc_set_batch_invariant": [c_int],
    "lmsfc_xchg_init": [],
    "lmsfc_xchg_reduce": [POINTER(XchgPeers), c_int, c_int64, c_int64, P, c_int, P],
    "lmsfc_xchg_norm_publish": [POINTER(XchgPeers), P, c_int, c_int, P, P],
    "lmsfc_xchg_norm_wait": [POINTER(XchgPeers), c_int, P, P],
    "lmsfc_xchg_adam_push": [POINTER(XchgPeers), c_int, c_int64, c_int64, c_int64, c_int64, c_int64, P, P, c_float, c_float,
                             c_float, P, c_int, P],
    "lmsfc_xchg_wait": [POINTER(XchgPeers), c_int, P],
    "lmsfc_gc_build_indexes": [P, c_int64, P, c_int, c_float, P, P],
    "lmsfc_step_capture_begin": [P],
    "lmsfc_step_capture_end": [P, POINTER(c_void_p), POINTER(c_int)],
    "lmsfc_step_launch": [P, P],
    "lmsfc_step_destroy": [P],
}
EXPORTS = ["lmsfc_last_error", "lmsfc_version", "lmsfc_launch_count", "lmsfc_rans_lane_cap",
           "lmsfc_rans_checkpoints"] + list(_SIGS)

lib.lmsfc_last_error.restype = c_char_p
lib.lmsfc_last_error.argtypes = []
lib.lmsfc_version.restype = c_int
lib.lmsfc_version.argtypes = []
lib.lmsfc_launch_count.restype = c_uint64
lib.lmsfc_launch_count.argtypes = []
lib.lmsfc_rans_lane_cap.restype = c_int64
lib.lmsfc_rans_lane_cap.argtypes = [c_int64, c_int]
lib.lmsfc_rans_checkpoints.restype = c_int
lib.lmsfc_rans_checkpoints.argtypes = [c_int64, c_int, c_int]


def launch_count():
    return int(lib.lmsfc_launch_count())


def _wrap(name, argtypes):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = c_int

    def call(*args):
        rc = fn(*args)
        if rc != 0:
            msg = lib.lmsfc_last_error().decode("utf-8", "replace")
            if rc in (1, 3):
                raise ValueError(f"{name}: {msg}")
            raise RuntimeError(f"{name}: {msg}")
    call.__name__ = name
    return call


for _n, _a in _SIGS.items():
    globals()[_n[len("lmsfc_"):]] = _wrap(_n, _a)


_FUSABLE = {}
# Where the GDN contraction is fused into the producing GEMM in TRAINING (bit mask; LMSFC_FUSE overrides, for A/B runs):
# 1 conv/deconv forward + GDN; 2 data gradient + GDN backward; 4 also where the GDN output has a second consumer whose
# gradient must be added first (the synthesis transform's heads: serialises head.dgrad -> up.dgrad).
# Measured in the captured batch-16 step on a B200 (profiles/r2_fuse_ab.txt; ms/step, +-0.05 %): none 1.000, 1: 1.015,
# 3: 1.007, 7: 1.014, 2: 0.966-0.990, 6: 0.999.  Solo, every fused kernel beats its separate launches (profiles/
# r2_fused_bench.txt: conv2+gdn2 57.2 vs 60.5 us, conv3.dgrad+gdn2 32.1 vs 38.6 us) and halves the layer's HBM traffic;
# inside the step the forward fusion loses: the separate square / GDN kernels overlapped the convolution's tail through
# programmatic dependent launch, the fused CTA (512 TMEM columns, 213 KB) keeps its SM to itself until it exits.
# Inference (no gradient) always takes the fused forward: y is the only tensor written.
# Re-measured on the final round-2 tree (profiles/tools/fuse_ab2.sh, same box, two runs each): 2: 0.9661 / 0.9673,
# 6: 0.9643 / 0.9649, 7: 0.9835 / 0.9825, 0: 0.972, 3: 0.986 -> 6 (the heads' gradient rides into the fused kernel as addend).
FUSE_MODE = int(os.environ.get("LMSFC_FUSE", "6"))


def gdn_fusable(desc, dgrad=False, n_in=0, inference=False):
    """True when conv `desc` (forward, or its data gradient w.r.t. the first n_in channels) can carry a fused GDN."""
    if not inference and not (FUSE_MODE & (2 if dgrad else 1)):
        return False
    key = (tuple(getattr(desc, f) for f, _ in ConvDesc._fields_), bool(dgrad), int(n_in))
    r = _FUSABLE.get(key)
    if r is None:
        out = c_int(0)
        conv_gdn_fusable(desc, int(dgrad), int(n_in), ctypes.byref(out))
        r = _FUSABLE[key] = bool(out.value)
    return r


class batch_invariant:
    """Context manager around lmsfc_set_batch_invariant (the entropy-coded path's h_s)."""

    def __enter__(self):
        set_batch_invariant(1)
        return self

    def __exit__(self, *a):
        set_batch_invariant(0)
        return False


def ptr(t):
    """Device (or host) address of a tensor, None -> NULL."""
    return None if t is None else t.data_ptr()


def stream():
    import torch
    return torch.cuda.current_stream().cuda_stream
