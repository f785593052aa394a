"""ctypes signatures of include/pdb200.h."""
import ctypes
from ctypes import POINTER, Structure, c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_void_p

P = c_void_p


class SyncPeersC(Structure):
    """pdb_sync_peers of include/pdb200.h."""
    _fields_ = [("staging", c_void_p * 16), ("pads", c_void_p * 16), ("rank", c_int), ("world", c_int),
                ("slot_stride", c_int), ("row", c_int)]


class ConvShape(Structure):
    _fields_ = [("N", c_int), ("H", c_int), ("W", c_int), ("C", c_int), ("K", c_int), ("R", c_int), ("S", c_int),
                ("stride", c_int), ("pad", c_int)]


def declare(L):
    def sig(name, args, res=c_int):
        f = getattr(L, name)
        f.argtypes, f.restype = args, res

    sig("pdb_last_error", [], c_char_p)
    sig("pdb_version", [])
    sig("pdb_sgd_step", [P, P, P, P, c_int, c_int64, c_float, c_float, c_float, c_float, c_int, c_int, c_int, P, P,
                         c_int, P])
    sig("pdb_check_finite", [P, c_int64, P, P])
    sig("pdb_scaler_update", [P, P, P, P, c_float, c_float, c_int, P])
    sig("pdb_cast_f32_to_16", [P, P, c_int, c_int64, P])
    sig("pdb_xent_fwd_bwd", [P, c_int, P, c_int, c_int, P, P, c_int, P, P])
    sig("pdb_topk_correct", [P, c_int, P, c_int, c_int, P, c_int, P, P])
    sig("pdb_bn_stats_scratch_bytes", [c_int], c_size_t)
    sig("pdb_bn_stats", [P, c_int, c_int64, c_int, P, P, c_size_t, P])
    sig("pdb_bn_apply", [P, P, P, c_int, c_int64, c_int, P, c_double, P, P, c_float, c_float, P, P, P, P, P, c_int, P, P])
    sig("pdb_bn_apply_eval", [P, P, P, c_int, c_int64, c_int, P, P, P, P, c_float, c_int, P])
    sig("pdb_bn_bwd_reduce", [P, P, P, P, c_int, c_int64, c_int, P, P, P, P, POINTER(SyncPeersC), P])
    sig("pdb_bn_bwd_reduce_scratch_floats", [c_int], c_size_t)
    sig("pdb_bn_bwd_reduce_det", [P, P, P, P, c_int, c_int64, c_int, P, P, P, P, c_size_t, P])
    sig("pdb_bn_bwd_apply", [P, P, P, P, P, P, c_int, c_int64, c_int, P, P, P, P, c_double, P, P, P])
    sig("pdb_maxpool3x3s2_fwd", [P, P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_maxpool3x3s2_bwd", [P, P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_avgpool_fwd", [P, P, c_int, c_int, c_int, c_int, P])
    sig("pdb_avgpool_bwd", [P, P, c_int, c_int, c_int, c_int, P])
    sig("pdb_nchw_f32_to_nhwc16", [P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_stem_im2col", [P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_stem_s2d", [P, P, c_int, c_int, c_int, c_int, P])
    sig("pdb_stem_weight_s2d", [P, P, c_int, c_int, P])
    sig("pdb_stem_fprop", [P, P, P, c_int, c_int, c_int, c_int, P, P, POINTER(SyncPeersC), P])
    sig("pdb_stem_wgrad", [P, P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_stem_dw_unpack", [P, P, c_int, c_int, P])
    sig("pdb_add16", [P, P, P, c_int, c_int64, P])
    sig("pdb_upsample2_zero", [P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_add_strided2", [P, P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_conv_fprop", [P, P, P, c_int, POINTER(ConvShape), c_int64, c_int64, P, P, P, c_int, c_int, POINTER(SyncPeersC), P])
    sig("pdb_conv_stats_scratch_floats", [c_int], c_size_t)
    sig("pdb_conv_wgrad", [P, P, P, c_int, POINTER(ConvShape), c_int64, c_int64, c_int64, c_int64, c_int, c_int, P])
    sig("pdb_weight_dgrad_layout", [P, P, c_int, c_int, c_int, c_int, c_int, P])
    sig("pdb_weight_dgrad_layout_batched", [P, P, P, c_int, c_int, P])
    sig("pdb_pad_rows16", [P, P, c_int, c_int, c_int, c_int, P])
    sig("pdb_colsum16", [P, P, c_int, c_int, c_int, P])
    for name, args in OPTIONAL.items():
        if hasattr(L, name):
            sig(name, args)


OPTIONAL = {
    "pdb_allreduce_inplace": [P, P, P, P, c_int, c_int, c_int64, c_int64, c_float, P, c_int, P],
    "pdb_allreduce_inplace_wire16": [P, P, P, P, c_int64, c_int, c_int, c_int, c_int64, c_int64, c_float, P, c_int, P],
    "pdb_bn_stats_sync": [P, c_int, c_int64, c_int, P, P, c_size_t, P, P, c_int, c_int, c_int, P],
    "pdb_broadcast_inplace": [P, P, c_int, c_int, c_int, c_int64, c_int64, c_int, P],
    "pdb_small_allreduce": [P, P, c_int, c_int, P, c_int, c_int, P],
    "pdb_allreduce_oneshot": [P, P, P, c_int, c_int, c_int64, c_int64, c_float, P, P],
    "pdb_peer_pull": [P, P, c_int64, P],
    "pdb_reduce_to_root": [P, P, c_int, c_int64, P],
    "pdb_peer_gather": [P, P, c_int, c_int64, c_int64, c_int, P],
    "pdb_comm_max_blocks": [c_int, c_size_t],
    "pdb_comm_set_trace": [P, c_int],
}
