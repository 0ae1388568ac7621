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

Strided convolutions over thin inputs (AlexNet conv1: RGB, 11x11, stride 4) run as ``S2dConvPlan``: a space-to-depth
copy of the input turns them into a stride-1 convolution over s*s*C channels (3x3 over 48 channels for conv1), which
has pixels wide enough for full tile loads and an implicit weight gradient; ``make_plan`` chooses.
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
    def forward(self, x, y2d, bias=None, act=ACT_NONE, wait=None, tile=None):
        """y2d [P][F] (a view of the channels-last output) = act(conv(x) + bias)."""
        self._x_raw = x
        self._x_saved = self._x_operand(x)
        K.conv_rows(self._x_saved, self.g_fwd, self.wf, self.F, y2d, bias=bias, act=act, wait=wait, tile=tile)

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

    def dgrad(self, dy, dx2d, grad_act=ACT_NONE, aux2d=None, tile=None):
        """dx2d [N*H*W][C] = conv^T(dY) (times act'(aux) when the source activation's backward is fused here)."""
        dyo = self._dy_operand(dy, 1)
        if self.dgrad_implicit:
            K.conv_rows(dyo, self.g_dgrad, self.wd, self.C, dx2d, grad_act=grad_act, aux=aux2d, tile=tile)
            return
        kc = self.kk * self.C
        dy2d = _mat(dyo, self.P, self.F, self.ldf)
        # dcol[P][(tap,c)] = dY[P][F] . Wdense[F][(tap,c)]
        out = {"out32": self.dcol[:, :kc]} if self.dtype == torch.float32 else {"out16": self.dcol[:, :kc]}
        K.gemm(dy2d, self.w_dense[:, :kc], self.P, kc, self.F, b_mn=True, **out)
        K.col2im_nhwc(self.dcol, self.dcol.stride(0), dx2d, self.N, self.H, self.W, self.C, self.OH, self.OW, self.k,
                      self.stride, self.pad, grad_act, aux2d)


def s2d_filter_index(C: int, k: int, s: int, Cs: int) -> torch.Tensor:
    """idx[j] = column of the reference filter layout [F][(c*k + ky)*k + kx] that column j of the space-to-depth filter
    [F][(cs*ks + ty)*ks + tx] takes its value from (cs = (dy*s + dx)*C + c, ky = ty*s + dy, kx = tx*s + dx), or -1 for
    taps beyond k and for the padding channels cs >= s*s*C."""
    ks = (k + s - 1) // s
    idx = torch.full((Cs * ks * ks,), -1, dtype=torch.int32)
    for dy in range(s):
        for dx in range(s):
            for c in range(C):
                cs = (dy * s + dx) * C + c
                for ty in range(ks):
                    for tx in range(ks):
                        ky, kx = ty * s + dy, tx * s + dx
                        if ky < k and kx < k:
                            idx[(cs * ks + ty) * ks + tx] = (c * k + ky) * k + kx
    return idx


def space_to_depth_ref(x: torch.Tensor, s: int, pad: int, Hs: int, Ws: int, Cs: int) -> torch.Tensor:
    """Plain-torch statement of ``sg_space_to_depth``: x [N,C,H,W] -> [N,Cs,Hs,Ws] with channel (dy*s + dx)*C + c of
    pixel (Y, X) = x[n][c][Y*s + dy - pad][X*s + dx - pad] (zero outside)."""
    N, C, H, W = x.shape
    xp = torch.zeros(N, C, Hs * s, Ws * s, dtype=x.dtype, device=x.device)
    h = min(H, Hs * s - pad)
    w = min(W, Ws * s - pad)
    xp[:, :, pad:pad + h, pad:pad + w] = x[:, :, :h, :w]
    out = torch.zeros(N, Cs, Hs, Ws, dtype=x.dtype, device=x.device)
    for dy in range(s):
        for dx in range(s):
            out[:, (dy * s + dx) * C:(dy * s + dx + 1) * C] = xp[:, :, dy::s, dx::s]
    return out


class S2dConvPlan:
    """k x k, stride s > 1 over C thin channels == ceil(k/s)^2, stride 1 over s*s*C channels of the space-to-depth
    image (taps beyond k are zero filters).  Same interface as ``ConvPlan`` for an input layer (no data gradient).

    x[n][Y*s+dy-pad][X*s+dx-pad][c] -> xs[n][Y][X][(dy*s+dx)*C + c];  filter tap (ty*s+dy, tx*s+dx) of channel c ->
    inner tap (ty, tx) of channel (dy*s+dx)*C + c.  The inner plan keeps the reference's [F][c*k*k + tap] layout
    for its own (inner) geometry, so the re-layouts are two index lists over the columns of the weight matrix."""

    def __init__(self, N, C, H, W, F, k, stride, pad, dtype: torch.dtype, device):
        assert stride > 1
        self.N, self.C, self.H, self.W, self.F, self.k, self.stride, self.pad = N, C, H, W, F, k, stride, pad
        self.dtype, self.device = dtype, device
        eb = 2 if dtype == torch.bfloat16 else 4
        q = 16 // eb
        s = stride
        self.ks = (k + s - 1) // s
        self.OH = (H + 2 * pad - k) // s + 1
        self.OW = (W + 2 * pad - k) // s + 1
        self.P = N * self.OH * self.OW
        self.Hs, self.Ws = self.OH + self.ks - 1, self.OW + self.ks - 1
        self.Cs = _rup(s * s * C, q)
        self.inner = ConvPlan(N, self.Cs, self.Hs, self.Ws, F, self.ks, 1, 0, dtype, device, need_dgrad=False)
        self.wf = self.inner.wf
        self.xs = torch.zeros(N, self.Hs, self.Ws, self.Cs, dtype=dtype, device=device)
        self.xs_cl = self.xs.permute(0, 3, 1, 2)          # [N, Cs, Hs, Ws] channels-last view for the inner plan
        ncol_in = self.Cs * self.ks * self.ks
        self.idx = s2d_filter_index(C, k, s, self.Cs).to(device)
        self.w_inner = torch.zeros(F, ncol_in, dtype=torch.float32, device=device)
        self.dw_inner = torch.zeros(F, ncol_in, dtype=torch.float32, device=device)

    def pack(self, w: torch.Tensor) -> None:
        K.gather_cols(w, self.w_inner, self.idx)
        self.inner.pack(self.w_inner)

    def forward(self, x, y2d, bias=None, act=ACT_NONE, wait=None):
        K.space_to_depth(x, self.xs, self.N, self.H, self.W, self.C, self.Hs, self.Ws, self.Cs, self.stride, self.pad)
        self.inner.forward(self.xs_cl, y2d, bias=bias, act=act, wait=wait)

    def wgrad(self, dy, dw, accumulate=False):
        self.inner.wgrad(dy, self.dw_inner)
        K.scatter_cols(self.dw_inner, dw, self.idx, accumulate)

    def dgrad(self, dy, dx2d, grad_act=ACT_NONE, aux2d=None):
        raise NotImplementedError("S2dConvPlan serves input layers only")


def make_plan(N, C, H, W, F, k, stride, pad, dtype: torch.dtype, device, need_dgrad: bool = True):
    """The plan for one convolution layer: space-to-depth for strided convolutions whose input pixels are narrower
    than 64 bytes and that need no data gradient (input layers), the direct implicit GEMM otherwise."""
    eb = 2 if dtype == torch.bfloat16 else 4
    if stride > 1 and not need_dgrad and C * eb < 64 and stride * stride * C * eb >= 64:
        return S2dConvPlan(N, C, H, W, F, k, stride, pad, dtype, device)
    return ConvPlan(N, C, H, W, F, k, stride, pad, dtype, device, need_dgrad=need_dgrad)
