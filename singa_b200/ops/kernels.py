"""Thin Python wrappers over the native sm_100a kernels (one function per entry point of
``include/singa_b200.h``).  They take CUDA torch tensors, launch on the current torch stream and write their
outputs in place.  4-D activations are channels-last (N,H,W,C contiguous)."""
from __future__ import annotations

import ctypes

import torch

from . import native
from ._decl import ShardTable, Updater
from .gemm import ACT_NONE, ACT_RELU, ACT_STANH, gemm  # noqa: F401  (re-exported)


def _lib():
    return native.require()


def _bf(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return 1
    assert t.dtype == torch.float32, t.dtype
    return 0


def _call(name: str, *args):
    rc = getattr(_lib(), name)(*args, native.stream_ptr())
    native.check(rc, name)
    native.count_launch()


def act_fwd(x, y, act):
    _call("sg_act_fwd", x.data_ptr(), y.data_ptr(), x.numel(), act, _bf(x))


def act_bwd(dy, y, dx, act):
    _call("sg_act_bwd", dy.data_ptr(), y.data_ptr(), dx.data_ptr(), dy.numel(), act, _bf(dy))


def dropout(x, y, pkeep: float, seed: int, seed_dev=None):
    _call("sg_dropout", x.data_ptr(), y.data_ptr(), x.numel(), float(pkeep), seed & (2**64 - 1),
          native.ptr(seed_dev), _bf(x))


def colsum(x, rows, cols, ld, out, accumulate=False):
    _call("sg_colsum", x.data_ptr(), ld, rows, cols, out.data_ptr(), 1 if accumulate else 0, _bf(x))


def cast_pack(src, lds, dst, ldd, rows, cols, perm=(0, 0)):
    _call("sg_cast_pack", src.data_ptr(), lds, dst.data_ptr(), ldd, rows, cols, perm[0], perm[1])


def repack_bf16(src, lds, dst, ldd, rows, cols, perm=(0, 0)):
    _call("sg_repack_bf16", src.data_ptr(), lds, dst.data_ptr(), ldd, rows, cols, perm[0], perm[1])


def convert(src, dst):
    _call("sg_convert", src.data_ptr(), _bf(src), dst.data_ptr(), _bf(dst), src.numel())


def copy2d(src, lds, dst, ldd, rows, cols):
    assert src.dtype == dst.dtype
    _call("sg_copy2d", src.data_ptr(), lds, dst.data_ptr(), ldd, rows, cols, _bf(src))


copy2d_bf16 = copy2d


IP_SMALL_MAX_N = 16


def ip_small_fwd(x, w, y, M, K, N, bias=None, act=ACT_NONE):
    """y[M][N] = act(x[M][K] . w[K][N] + bias), N <= 16 (csrc/ip_small.cu)."""
    assert x.stride(-1) == 1 and w.stride(-1) == 1 and y.stride(-1) == 1
    _call("sg_ip_small_fwd", x.data_ptr(), x.stride(0), _bf(x), w.data_ptr(), w.stride(0), _bf(w), native.ptr(bias),
          _bf(bias) if bias is not None else 0, y.data_ptr(), y.stride(0), _bf(y), M, K, N, act)


def ip_small_dgrad(dy, w, dx, M, K, N, grad_act=ACT_NONE, aux=None):
    """dx[M][K] = (dy[M][N] . w[K][N]^T) * act'(aux)."""
    assert dy.stride(-1) == 1 and w.stride(-1) == 1 and dx.stride(-1) == 1
    _call("sg_ip_small_dgrad", dy.data_ptr(), dy.stride(0), _bf(dy), w.data_ptr(), w.stride(0), _bf(w), native.ptr(aux),
          aux.stride(0) if aux is not None else 0, _bf(aux) if aux is not None else 0, grad_act, dx.data_ptr(),
          dx.stride(0), _bf(dx), M, K, N)


def ip_small_wgrad_workspace(M, K, device):
    """Zeroed workspace for ``ip_small_wgrad`` (one per layer: calls that may overlap must not share it)."""
    lib = _lib()
    lib.sg_ip_small_wgrad_ws_bytes.restype = ctypes.c_longlong
    nbytes = int(lib.sg_ip_small_wgrad_ws_bytes(int(M), int(K)))
    return torch.zeros((nbytes + 3) // 4, dtype=torch.float32, device=device)


def ip_small_wgrad(x, dy, dw, M, K, N, ws, db=None, accumulate=False):
    """dw[K][N] (+)= x[M][K]^T . dy[M][N]; db[N] (+)= column sums of dy.  Deterministic (no atomics on the result)."""
    assert dw.dtype == torch.float32 and (db is None or db.dtype == torch.float32) and ws.dtype == torch.float32
    _call("sg_ip_small_wgrad", x.data_ptr(), x.stride(0), _bf(x), dy.data_ptr(), dy.stride(0), _bf(dy), dw.data_ptr(),
          dw.stride(0), native.ptr(db), M, K, N, 1 if accumulate else 0, ws.data_ptr())


def space_to_depth(x, xs, N, H, W, C, Hs, Ws, Cs, s, pad):
    assert x.dtype == xs.dtype
    _call("sg_space_to_depth", x.data_ptr(), xs.data_ptr(), N, H, W, C, Hs, Ws, Cs, s, pad, _bf(x))


def gather_cols(src, dst, idx):
    """dst[r][j] = src[r][idx[j]] (0 where idx[j] < 0); src fp32 / bf16, dst fp32."""
    assert dst.dtype == torch.float32 and idx.dtype == torch.int32 and src.stride(-1) == 1 and dst.stride(-1) == 1
    _call("sg_gather_cols", src.data_ptr(), _bf(src), src.stride(0), dst.data_ptr(), dst.stride(0), dst.shape[0],
          idx.numel(), idx.data_ptr())


def scatter_cols(src, dst, idx, accumulate=False):
    """dst[r][idx[j]] (+)= src[r][j] for idx[j] >= 0; fp32 both."""
    assert src.dtype == torch.float32 and dst.dtype == torch.float32 and idx.dtype == torch.int32
    _call("sg_scatter_cols", src.data_ptr(), src.stride(0), dst.data_ptr(), dst.stride(0), src.shape[0], idx.numel(),
          idx.data_ptr(), 1 if accumulate else 0)


def transpose_planes(src, dst, batch, rows, cols):
    _call("sg_transpose_planes", src.data_ptr(), _bf(src), dst.data_ptr(), _bf(dst), batch, rows, cols)


def asum(x, out):
    _call("sg_asum", x.data_ptr(), x.numel(), out.data_ptr(), _bf(x))


def fill_random(dst, kind: int, a: float, b: float, scale: float, seed: int):
    _call("sg_fill_random", dst.data_ptr(), dst.numel(), kind, float(a), float(b), float(scale),
          seed & (2**64 - 1))


def _src_f32(t: torch.Tensor) -> int:
    assert t.dtype in (torch.uint8, torch.float32), t.dtype
    return 1 if t.dtype == torch.float32 else 0


def mnist_parse(src, dst, norm_a, norm_b):
    _call("sg_mnist_parse_src", src.data_ptr(), _src_f32(src), dst.data_ptr(), src.numel(), float(norm_a),
          float(norm_b), _bf(dst))


def mnist_distort(src, dst, n_img, in_size, out_size, affine, disp, norm_a, norm_b):
    _call("sg_mnist_distort", src.data_ptr(), _src_f32(src), dst.data_ptr(), n_img, in_size, out_size,
          affine.data_ptr(), native.ptr(disp), float(norm_a), float(norm_b), _bf(dst))


def rgb_parse(src, mean, dst, batch, C, H, W, crop, off_h, off_w, mirror, scale):
    _call("sg_rgb_parse_src", src.data_ptr(), _src_f32(src), native.ptr(mean), dst.data_ptr(), batch, C, H, W, crop,
          native.ptr(off_h), native.ptr(off_w), native.ptr(mirror), float(scale), _bf(dst))


def softmax_loss(logits, ldl, labels, rows, dim, topk, scale, prob, grad, ldg, metric):
    _call("sg_softmax_loss", logits.data_ptr(), ldl, labels.data_ptr(), rows, dim, topk, float(scale),
          native.ptr(prob), prob.stride(0) if prob is not None else 0, native.ptr(grad), ldg,
          _bf(grad) if grad is not None else 0, native.ptr(metric))


def pool_fwd(x, y, N, H, W, C, OH, OW, k, s, pad, is_max):
    _call("sg_pool_fwd", x.data_ptr(), y.data_ptr(), N, H, W, C, OH, OW, k, s, pad, 1 if is_max else 0, _bf(x))


def pool_bwd(x, y, dy, dx, N, H, W, C, OH, OW, k, s, pad, is_max):
    _call("sg_pool_bwd", x.data_ptr(), y.data_ptr(), dy.data_ptr(), dx.data_ptr(), N, H, W, C, OH, OW, k, s,
          pad, 1 if is_max else 0, _bf(x))


def lrn_fwd(x, y, norm, pixels, C, lsize, alpha, beta, knorm):
    _call("sg_lrn_fwd", x.data_ptr(), y.data_ptr(), norm.data_ptr(), pixels, C, lsize, float(alpha),
          float(beta), float(knorm), _bf(x))


def lrn_bwd(x, norm, dy, dx, pixels, C, lsize, alpha, beta):
    _call("sg_lrn_bwd", x.data_ptr(), norm.data_ptr(), dy.data_ptr(), dx.data_ptr(), pixels, C, lsize,
          float(alpha), float(beta), _bf(x))


def im2col_nhwc(x, col, ld, N, H, W, C, OH, OW, k, s, pad):
    _call("sg_im2col", x.data_ptr(), col.data_ptr(), ld, N, H, W, C, OH, OW, k, s, pad, _bf(x))


def col2im_nhwc(dcol, ld, dx, N, H, W, C, OH, OW, k, s, pad, grad_act=ACT_NONE, aux=None):
    _call("sg_col2im", dcol.data_ptr(), ld, dx.data_ptr(), N, H, W, C, OH, OW, k, s, pad, grad_act,
          native.ptr(aux), _bf(dx))


# ---------------------------------------------------------------------------------------------------------
# Implicit-GEMM convolution (csrc/conv_implicit.cu)
# ---------------------------------------------------------------------------------------------------------
def conv_chunk_bytes(channels: int, elem_bytes: int) -> int:
    return _lib().sg_conv_chunk_bytes(channels, elem_bytes)


def conv_geom(N, H, W, C, ldc, OH, OW, k, stride, pad_lo, chunk_bytes=128):
    from ._decl import ConvGeom
    return ConvGeom(N, H, W, C, ldc, OH, OW, k, stride, pad_lo, chunk_bytes)


def _epilogue(out, bias=None, act=ACT_NONE, grad_act=ACT_NONE, aux=None, wait=None):
    ep = native.GemmEpilogue()
    assert out.stride(-1) == 1 or out.shape[-1] == 1
    if out.dtype == torch.float32:
        ep.d32, ep.ldd32 = out.data_ptr(), out.stride(0)
    else:
        assert out.dtype == torch.bfloat16
        ep.d16, ep.ldd16 = out.data_ptr(), out.stride(0)
    if bias is not None:
        ep.bias = bias.data_ptr()
        ep.bias_bf16 = _bf(bias)
    ep.act, ep.grad_act = act, grad_act
    if grad_act != ACT_NONE:
        assert aux is not None and (aux.stride(-1) == 1 or aux.shape[-1] == 1)
        ep.aux, ep.ldaux, ep.aux_bf16 = aux.data_ptr(), aux.stride(0), _bf(aux)
    if wait is not None:
        flags, seq = wait
        ep.wait_flags, ep.nwait, ep.wait_seq = flags.data_ptr(), flags.numel(), seq
    return ep


def conv_rows(src, geom, filt, ncols, out2d, *, bias=None, act=ACT_NONE, grad_act=ACT_NONE, aux=None, wait=None,
              tile=None):
    """out2d[pixel][n] = epilogue(sum_k patch(src; pixel, k) * filt[n][k]) — forward and stride-1 data gradient.
    ``tile`` = (bn, pair) forces the tile width (0 = library's choice) and single-CTA (1) / CTA-pair (2) tiles."""
    ep = _epilogue(out2d, bias, act, grad_act, aux, wait)
    if tile is not None:
        ep.tile_bn, ep.tile_pair = tile
    assert filt.dtype == src.dtype and filt.stride(-1) == 1
    rc = _lib().sg_conv_rows(0 if _bf(src) else 1, src.data_ptr(), ctypes.byref(geom), filt.data_ptr(), filt.stride(0),
                             ncols, ctypes.byref(ep), native.stream_ptr())
    native.check(rc, "sg_conv_rows")
    native.count_launch()


def conv_wgrad(src, geom, dy2d, F, dw, accumulate=False):
    """dw[f][c*k*k + tap] (+)= sum_pixel dy2d[pixel][f] * patch(src; pixel, tap, c)."""
    assert dy2d.dtype == src.dtype and dy2d.stride(-1) == 1 and dw.dtype == torch.float32 and dw.stride(-1) == 1
    rc = _lib().sg_conv_wgrad(0 if _bf(src) else 1, src.data_ptr(), ctypes.byref(geom), dy2d.data_ptr(), dy2d.stride(0),
                              F, dw.data_ptr(), dw.stride(0), 1 if accumulate else 0, native.stream_ptr())
    native.check(rc, "sg_conv_wgrad")
    native.count_launch(1 if accumulate else 2)


def conv_pack(w, ldw, F, C, k, fwd, cp_f, dgrad=None, fp_d=0):
    _call("sg_conv_pack", w.data_ptr(), _bf(w), ldw, F, C, k, fwd.data_ptr(), fwd.stride(0), cp_f,
          native.ptr(dgrad), dgrad.stride(0) if dgrad is not None else 0, fp_d, _bf(fwd))


UPD_SGD, UPD_NESTEROV, UPD_ADAGRAD, UPD_RMSPROP, UPD_ADADELTA = range(5)
ARENA_CHUNK = 1024


def update_arena(w, g, h, u, w16, chunk_param, lr_mult, wd_mult, first_chunk, nchunks, *, kind, lr,
                 momentum=0.0, weight_decay=0.0, delta=1e-7, rho=0.9, grad_scale=1.0, lr_dev=None):
    up = Updater(kind, lr, momentum, weight_decay, delta, rho, grad_scale, native.ptr(lr_dev))
    _call("sg_update_arena", w.data_ptr(), g.data_ptr(), native.ptr(h), native.ptr(u), native.ptr(w16),
          chunk_param.data_ptr(), lr_mult.data_ptr(), wd_mult.data_ptr(), first_chunk, nchunks,
          ctypes.byref(up))


def easgd_sync(w, server, alpha, w16=None):
    _call("sg_easgd_sync", w.data_ptr(), server.data_ptr(), w.numel(), float(alpha), native.ptr(w16))


def randomsync(w, snap, server, ratio, seed, w16=None, nselected=None):
    _call("sg_randomsync", w.data_ptr(), snap.data_ptr(), server.data_ptr(), w.numel(), float(ratio),
          seed & (2**64 - 1), native.ptr(w16), native.ptr(nselected))


def shard_table(ptrs, per: int) -> ShardTable:
    t = ShardTable()
    for k, p in enumerate(ptrs):
        t.server[k] = p
    t.nshards, t.per = len(ptrs), per
    return t


def easgd_sync_sharded(w, table: ShardTable, alpha, w16=None):
    _call("sg_easgd_sync_sharded", w.data_ptr(), ctypes.byref(table), w.numel(), float(alpha), native.ptr(w16))


def randomsync_sharded(w, snap, table: ShardTable, ratio, seed, seed_dev=None, w16=None, nselected=None):
    _call("sg_randomsync_sharded", w.data_ptr(), snap.data_ptr(), ctypes.byref(table), w.numel(), float(ratio),
          seed & (2**64 - 1), native.ptr(seed_dev), native.ptr(w16), native.ptr(nselected))


def group_average(mine, table: ShardTable, lo: int, n: int):
    _call("sg_group_average", mine.data_ptr(), ctypes.byref(table), lo, n)
