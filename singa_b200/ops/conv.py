"""Convolution as implicit GEMM on the tcgen05 kernels (``csrc/conv_implicit.cu``): geometry, operand buffers and
the three launches (forward, data gradient, weight gradient) of one convolution layer.

Reference op: ConvolutionLayer, src/worker/layer.cc:18-123 — per image ``unpack_patch2col`` + ``dot`` forward,
``dot`` + ``pack_col2patch`` backward, weight ``{F, C*k*k}`` with column ``(c*k + ky)*k + kx``.

Nothing column-shaped is written to HBM on the implicit paths: TMA im2col loads build the operand tiles inside the
GEMM.  What remains explicit (``sg_im2col`` / ``sg_col2im``), and why:
  * the WEIGHT gradient when a pixel of the input is narrower than 64 bytes (grey / RGB images): there the patch
    matrix is an MN-major operand, which for tf32 exists only with 128-byte rows — a 3-channel pixel would be 90 %
    zero fill;
  * the DATA gradient of strided convolutions (no such layer feeds a trainable layer in the five benchmark nets).
"""
from __future__ import annotations

import torch

from . import kernels as K
from .gemm import ACT_NONE


def _rup(n: int, m: int) -> int:
    return (n + m - 1) // m * m


def _mat(t: torch.Tensor, rows: int, cols: int, ld: int) -> torch.Tensor:
    """[rows][cols] matrix view (row stride ld) of a channels-last activation's memory."""
    return t.as_strided((rows, cols), (ld, 1), t.storage_offset())


class ConvPlan:
    def __init__(self, N, C, H, W, F, k, stride, pad, dtype: torch.dtype, device, need_dgrad: bool = True):
        assert dtype in (torch.float32, torch.bfloat16)
        self.N, self.C, self.H, self.W, self.F, self.k, self.stride, self.pad = N, C, H, W, F, k, stride, pad
        self.dtype, self.device = dtype, device
        eb = 2 if dtype == torch.bfloat16 else 4
        self.eb, self.q = eb, 16 // eb          # q: elements per 16 bytes (TMA stride granule)
        self.bk = 128 // eb                     # elements per 128-byte k-block
        self.kk = k * k
        self.OH = (H + 2 * pad - k) // stride + 1
        self.OW = (W + 2 * pad - k) // stride + 1
        self.P = N * self.OH * self.OW
        self.need_dgrad = need_dgrad
        z = lambda *shape: torch.zeros(*shape, dtype=dtype, device=device)   # noqa: E731

        # ---- forward operand: filters [F][(tap, c)] with Cp channels per tap
        self.chunk_f = K.conv_chunk_bytes(C, eb)
        ce = self.chunk_f // eb
        self.cp_f = _rup(C, ce)
        self.wf = z(F, _rup(self.kk * self.cp_f, self.bk))
        # the input must have 16-byte pixels: a padded copy when C is not a multiple of q
        self.ldc_in = _rup(C, self.q)
        self.x_pad = z(N, H, W, self.ldc_in) if self.ldc_in != C else None

        # ---- data gradient
        self.dgrad_implicit = stride == 1 and pad <= k - 1
        self.ldf = _rup(F, self.q)
        # dY copies with 16-byte pixels (one per consumer: the two gradients run on different streams)
        self.dy_pad = [z(N, self.OH, self.OW, self.ldf) for _ in range(2)] if self.ldf != F else None
        self.wd = None
        self.w_dense = None
        if need_dgrad:
            if self.dgrad_implicit:
                self.chunk_d = K.conv_chunk_bytes(F, eb)
                cd = self.chunk_d // eb
                self.fp_d = _rup(F, cd)
                self.wd = z(C, _rup(self.kk * self.fp_d, self.bk))
            else:
                self.w_dense = z(F, _rup(self.kk * C, self.q))     # [F][(tap, c)] dense, for the explicit dcol GEMM
                self.dcol = z(self.P, _rup(self.kk * C, self.q))

        # ---- weight gradient
        self.wgrad_implicit = C * eb >= 64
        self.col = None if self.wgrad_implicit else z(self.P, _rup(self.kk * C, self.q))

        self.g_fwd = K.conv_geom(N, H, W, C, self.ldc_in, self.OH, self.OW, k, stride, pad, self.chunk_f)
        self.g_wgrad = K.conv_geom(N, H, W, C, self.ldc_in, self.OH, self.OW, k, stride, pad, 128)
        if need_dgrad and self.dgrad_implicit:
            # rows = input pixels, source = dY: pixel (y, x) reads dY rows y - (k-1-pad) + j
            self.g_dgrad = K.conv_geom(N, self.OH, self.OW, F, self.ldf, H, W, k, 1, k - 1 - pad, self.chunk_d)

    # -- operands ------------------------------------------------------------------------------------------
    def pack(self, w: torch.Tensor) -> None:
        """Refresh the packed filters from ``w`` = [F][C*k*k] (fp32 master values or a mirror; any row stride)."""
        assert w.stride(-1) == 1
        K.conv_pack(w, w.stride(0), self.F, self.C, self.k, self.wf, self.cp_f,
                    self.wd if self.wd is not None else None, self.fp_d if self.wd is not None else 0)
        if self.w_dense is not None:
            K.conv_pack(w, w.stride(0), self.F, self.C, self.k, self.w_dense, self.C)

    def _x_operand(self, x: torch.Tensor) -> torch.Tensor:
        """``x``: channels-last activation ([N,C,H,W] channels_last, or [N,H,W] when C == 1)."""
        if self.x_pad is None:
            return x
        K.copy2d(x, self.C, self.x_pad, self.ldc_in, self.N * self.H * self.W, self.C)
        return self.x_pad

    def _dy_operand(self, dy: torch.Tensor, which: int) -> torch.Tensor:
        if self.dy_pad is None:
            return dy
        K.copy2d(dy, self.F, self.dy_pad[which], self.ldf, self.P, self.F)
        return self.dy_pad[which]

    # -- the three GEMMs -----------------------------------------------------------------------------------
    def forward(self, x, y2d, bias=None, act=ACT_NONE, wait=None):
        """y2d [P][F] (a view of the channels-last output) = act(conv(x) + bias)."""
        self._x_raw = x
        self._x_saved = self._x_operand(x)
        K.conv_rows(self._x_saved, self.g_fwd, self.wf, self.F, y2d, bias=bias, act=act, wait=wait)

    def wgrad(self, dy, dw, accumulate=False):
        """dw [F][C*k*k] (fp32, reference layout) (+)= dY^T patches(x saved by forward)."""
        dyo = self._dy_operand(dy, 0)
        dy2d = _mat(dyo, self.P, self.F, self.ldf)
        if self.wgrad_implicit:
            K.conv_wgrad(self._x_saved, self.g_wgrad, dy2d, self.F, dw, accumulate)
            return
        kc = self.kk * self.C
        K.im2col_nhwc(self._x_raw, self.col, self.col.stride(0), self.N, self.H, self.W, self.C, self.OH, self.OW,
                      self.k, self.stride, self.pad)
        K.gemm(dy2d, self.col[:, :kc], self.F, kc, self.P, a_mn=True, b_mn=True, out32=dw, accumulate=accumulate,
               perm=(self.C, self.kk) if self.kk > 1 and self.C > 1 else None)

    def dgrad(self, dy, dx2d, grad_act=ACT_NONE, aux2d=None):
        """dx2d [N*H*W][C] = conv^T(dY) (times act'(aux) when the source activation's backward is fused here)."""
        dyo = self._dy_operand(dy, 1)
        if self.dgrad_implicit:
            K.conv_rows(dyo, self.g_dgrad, self.wd, self.C, dx2d, grad_act=grad_act, aux=aux2d)
            return
        kc = self.kk * self.C
        dy2d = _mat(dyo, self.P, self.F, self.ldf)
        # dcol[P][(tap,c)] = dY[P][F] . Wdense[F][(tap,c)]
        out = {"out32": self.dcol[:, :kc]} if self.dtype == torch.float32 else {"out16": self.dcol[:, :kc]}
        K.gemm(dy2d, self.w_dense[:, :kc], self.P, kc, self.F, b_mn=True, **out)
        K.col2im_nhwc(self.dcol, self.dcol.stride(0), dx2d, self.N, self.H, self.W, self.C, self.OH, self.OW, self.k,
                      self.stride, self.pad, grad_act, aux2d)
