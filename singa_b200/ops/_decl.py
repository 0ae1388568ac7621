"""ctypes signatures of the non-GEMM entry points, in the order of include/singa_b200.h."""
import ctypes as C

vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong


class ConvGeom(C.Structure):
    _fields_ = [("N", i32), ("H", i32), ("W", i32), ("C", i32), ("ldc", i64), ("OH", i32), ("OW", i32),
                ("k", i32), ("stride", i32), ("pad_lo", i32), ("chunk_bytes", i32)]


class ShardTable(C.Structure):
    _fields_ = [("server", vp * 8), ("nshards", i32), ("per", i64)]


class Updater(C.Structure):
    _fields_ = [("kind", i32), ("lr", f32), ("momentum", f32), ("weight_decay", f32), ("delta", f32),
                ("rho", f32), ("grad_scale", f32), ("lr_dev", vp)]


SIGNATURES = {
    "sg_act_fwd": [vp, vp, i64, i32, i32, vp],
    "sg_act_bwd": [vp, vp, vp, i64, i32, i32, vp],
    "sg_dropout": [vp, vp, i64, f32, u64, vp, i32, vp],
    "sg_colsum": [vp, i64, i32, i32, vp, i32, i32, vp],
    "sg_cast_pack": [vp, i64, vp, i64, i32, i32, i32, i32, vp],
    "sg_repack_bf16": [vp, i64, vp, i64, i32, i32, i32, i32, vp],
    "sg_convert": [vp, i32, vp, i32, i64, vp],
    "sg_copy2d_bf16": [vp, i64, vp, i64, i32, i32, vp],
    "sg_copy2d": [vp, i64, vp, i64, i32, i32, i32, vp],
    "sg_transpose_planes": [vp, i32, vp, i32, i32, i32, i32, vp],
    "sg_ip_small_fwd": [vp, i64, i32, vp, i64, i32, vp, i32, vp, i64, i32, i32, i32, i32, i32, vp],
    "sg_ip_small_dgrad": [vp, i64, i32, vp, i64, i32, vp, i64, i32, i32, vp, i64, i32, i32, i32, i32, vp],
    "sg_ip_small_wgrad": [vp, i64, i32, vp, i64, i32, vp, i64, vp, i32, i32, i32, i32, vp, vp],
    "sg_space_to_depth": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "sg_gather_cols": [vp, i32, i64, vp, i64, i32, i32, vp, vp],
    "sg_scatter_cols": [vp, i64, vp, i64, i32, i32, vp, i32, vp],
    "sg_asum": [vp, i64, vp, i32, vp],
    "sg_fill_random": [vp, i64, i32, f32, f32, f32, u64, vp],
    "sg_mnist_parse": [vp, vp, i64, f32, f32, i32, vp],
    "sg_mnist_distort": [vp, i32, vp, i32, i32, i32, vp, vp, f32, f32, i32, vp],
    "sg_mnist_parse_src": [vp, i32, vp, i64, f32, f32, i32, vp],
    "sg_rgb_parse_src": [vp, i32, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, f32, i32, vp],
    "sg_rgb_parse": [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, f32, i32, vp],
    "sg_softmax_loss": [vp, i64, vp, i32, i32, i32, f32, vp, i64, vp, i64, i32, vp, vp],
    "sg_pool_fwd": [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "sg_pool_bwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "sg_lrn_fwd": [vp, vp, vp, i64, i32, i32, f32, f32, f32, i32, vp],
    "sg_lrn_bwd": [vp, vp, vp, vp, i64, i32, i32, f32, f32, i32, vp],
    "sg_im2col": [vp, vp, i64, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "sg_col2im": [vp, i64, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, i32, vp],
    "sg_conv_pack": [vp, i32, i64, i32, i32, i32, vp, i64, i32, vp, i64, i32, i32, vp],
    "sg_im2col_nhwc": [vp, vp, i64, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp],
    "sg_col2im_nhwc": [vp, i64, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, vp],
    "sg_update_arena": [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, C.POINTER(Updater), vp],
    "sg_easgd_sync": [vp, vp, i64, f32, vp, vp],
    "sg_randomsync": [vp, vp, vp, i64, f32, u64, vp, vp, vp],
    "sg_easgd_sync_sharded": [vp, C.POINTER(ShardTable), i64, f32, vp, vp],
    "sg_randomsync_sharded": [vp, vp, C.POINTER(ShardTable), i64, f32, u64, vp, vp, vp, vp],
    "sg_group_average": [vp, C.POINTER(ShardTable), i64, i64, vp],
    "sg_ps_push_update": [vp, vp],
    "sg_ps_push_update_noalloc": [vp, i32, vp],
    "sg_ps_flag_set": [vp, vp, i64, vp],
    "sg_ps_flag_wait": [vp, i32, vp, i64, vp, vp],
    "sg_ps_flag_bump": [vp, vp, vp],
    "sg_ipc_alloc": [i64, vp],
    "sg_ipc_free": [vp],
    "sg_ipc_get_handle": [vp, vp],
    "sg_ipc_open_handle": [vp, vp],
    "sg_ipc_close_handle": [vp],
}


def declare_all(lib):
    for name, args in SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = list(args)
            fn.restype = C.c_int
