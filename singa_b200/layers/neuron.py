"""Compute layers: Convolution, InnerProduct, Pooling, LRN, ReLU, Tanh, Dropout, SoftmaxLoss.

Reference: include/worker/layer.h:28-291, src/worker/layer.cc (citations per class).  Every layer has two
implementations selected once at setup: the native sm_100a path (tcgen05 GEMM + fused epilogues, bf16
channels-last activations) and the plain-PyTorch fp32 path of ``ops/reference.py`` (CPU oracle / library
baseline).  Forward/backward only launch kernels on pre-allocated buffers, so a whole training step can be
captured in a CUDA graph.

Fusions decided by ``NeuralNet.fuse_activations`` (natural fusion boundaries of SURVEY §2.5):
  * ``fused_act``      — a ReLU/Tanh that is the sole consumer of a conv / inner-product layer runs in that
                         layer's GEMM epilogue; the activation layer becomes an alias.
  * ``src_grad_act``   — when the source of a conv / inner-product layer is an activation layer, the data-gradient
                         GEMM (or col2im) multiplies by act'(y) in its epilogue and writes straight into the
                         activation's source gradient; the activation layer's backward is skipped.
"""
from __future__ import annotations

import os

import zlib

from typing import List, Optional

import torch

from ..device import ctx, pad8, step_state
from ..ops import reference as R
from ..ops.gemm import ACT_NONE, ACT_RELU, ACT_STANH
from ..param import param_factory
from ..proto import schema as S
from .base import Layer, alloc_act


def _K():
    from ..ops import kernels
    return kernels


class _Workspace:
    """One growing device scratch buffer shared by layers whose scratch use never overlaps in time
    (conv data-gradient column buffers, padded gradient copies)."""

    def __init__(self):
        self.buf: Optional[torch.Tensor] = None

    def reserve(self, nbytes: int) -> None:
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = torch.empty(nbytes, dtype=torch.uint8, device=ctx().device)

    def view(self, shape, dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.tensor([], dtype=dtype).element_size()
        assert self.buf is not None and self.buf.numel() >= nbytes, "workspace not reserved"
        return self.buf[:nbytes].view(dtype).view(shape)


workspace = _Workspace()


def _out(t: torch.Tensor) -> dict:
    """GEMM output keyword for an activation / gradient tensor of either compute dtype."""
    return {"out32": t} if t.dtype == torch.float32 else {"out16": t}


class NeuronLayer(Layer):
    """Shared fusion plumbing."""

    def _bias_operand(self):
        """Bias for the GEMM epilogue: the fp32 master value, or — when a sharded parameter server keeps the
        master copy elsewhere — this rank's bf16 mirror, which is what the servers broadcast into."""
        b = self.bias if self.bias.is_owner else self.bias.owner
        if getattr(b, "use_mirror", False):
            return b.mirror16
        return b.data

    def __init__(self):
        super().__init__()
        self.fused_act = ACT_NONE          # activation applied in this layer's forward epilogue
        self.src_grad_act = ACT_NONE       # activation gradient applied in this layer's dgrad epilogue
        self.src_grad_target: Optional[Layer] = None   # layer whose grad the dgrad writes (skips the act layer)

    def _src_grad(self):
        """(tensor to write the data gradient into, aux activation output or None)."""
        src = self.srclayers[0]
        if self.src_grad_act != ACT_NONE:
            return self.src_grad_target.grad_for(src), src.data_for(self)
        return src.grad_for(self), None

    def _need_src_grad(self) -> bool:
        src = self.srclayers[0]
        if src.is_datalayer or src.is_parserlayer:
            return False
        return src.grad_for(self) is not None or self.src_grad_act != ACT_NONE


# ================================================================================================
# InnerProduct
# ================================================================================================
class InnerProductLayer(NeuronLayer):
    """data = src . W + b with W {vdim, hdim}  (src/worker/layer.cc:162-213)."""
    type_id = "kInnerProduct"
    supports_fused_pull = True     # the forward GEMM can acquire the parameter server's version flags itself
    pull_wait = None

    def connection_type(self, k):
        return S.kOneToAll

    def setup(self):
        assert len(self.srclayers) == 1
        src_shape = self.srclayers[0].shape(self)
        self.batchsize = src_shape[0]
        self.vdim = 1
        for s in src_shape[1:]:
            self.vdim *= s
        self.hdim = int(self.proto.inner_product_param.num_output)
        if self.forced_shape is not None:   # one partition of a layer-partitioned layer
            self.hdim = self.forced_shape[1]
            self.proto.inner_product_param.num_output = self.hdim
        self.bias_term = self.proto.inner_product_param.bias_term
        self.feeds_loss = any(d.is_losslayer for d in self.dstlayers)
        self.shape_ = [self.batchsize, self.hdim]
        native = self.native
        self.data = alloc_act(self.shape_, torch.float32 if (self.feeds_loss and native) else None)
        self.grad = alloc_act(self.shape_)
        if not hasattr(self, "weight"):
            protos = list(self.proto.param)
            while len(protos) < 2:
                protos.append(S.ParamProto())
            self.weight = param_factory.create("Param")
            self.bias = param_factory.create("Param")
            if not protos[0].name:
                protos[0].name = "weight"
            if not protos[1].name:
                protos[1].name = "bias"
            # fan_in: the reference passes vdim*hdim (layer.cc:174, survey §2.8 item 4); the intent is vdim
            self.weight.setup(protos[0], [self.vdim, self.hdim], self.vdim)
            self.bias.setup(protos[1], [self.hdim], 0)
        self.src4d = len(src_shape) == 4 and src_shape[1] > 1 and src_shape[2] * src_shape[3] > 1
        if native:
            self._setup_native(src_shape)

    def _setup_native(self, src_shape):
        dev = ctx().device
        # flatten in the reference's (c,h,w) order: channels-last 4-D sources go through a transpose kernel;
        # a source whose row stride is not TMA-legal (multiple of 8) is re-strided into x_flat as well
        self.x_flat = None
        self.dx_flat = None
        self.ldq = 8 if ctx().is_bf16 else 4     # elements per 16 bytes: TMA row strides are multiples of this
        if self.src4d or self.vdim % self.ldq != 0:
            self.x_flat = alloc_act([self.batchsize, self.vdim])
            self.dx_flat = alloc_act([self.batchsize, self.vdim])
        # weight operand = the arena's mirror (or, fp32 on one GPU, the master values); when hdim is not a multiple
        # of 8 the arena stores the rows with a padded stride (ParamArena honours want_padded_ld): TMA-legal as is
        self.weight.want_padded_ld = True
        # classifier-sized layers: workspace of the deterministic weight-gradient kernel (allocated here, not inside a
        # step that may be under CUDA-graph capture)
        self._ips_ws = None
        if self.hdim <= 16 and self.batchsize >= 256 and dev.type == "cuda":
            self._ips_ws = _K().ip_small_wgrad_workspace(self.batchsize, self.vdim, dev)

    def get_params(self):
        return [self.weight, self.bias] if self.bias_term else [self.weight]

    def _flat2d(self, t):
        """2-D [batch, vdim] alias of ``t`` when its memory order already is the reference's flatten order."""
        if t.dim() == 2:
            return t
        v = t.reshape(self.batchsize, self.vdim)
        assert v.data_ptr() == t.data_ptr(), f"{self.name}: unexpected copy while flattening"
        return v

    def _x2d(self):
        src = self.srclayers[0].data_for(self)
        if not self.native:
            return src.reshape(self.batchsize, self.vdim).float()
        K = _K()
        if self.src4d:
            n, c, h, w = src.shape
            if self.vdim % self.ldq == 0:
                K.transpose_planes(src, self.x_flat, n, h * w, c)      # NHWC -> (c,h,w) order
            else:
                self.x_flat.copy_(src.reshape(self.batchsize, self.vdim))
            return self.x_flat
        x = self._flat2d(src)
        if x.stride(0) % self.ldq != 0:
            K.copy2d(x, x.stride(0), self.x_flat, self.x_flat.stride(0), self.batchsize, self.vdim)
            return self.x_flat
        return x

    def _w16(self):
        w = self.weight if self.weight.is_owner else self.weight.owner
        return w.data16

    def _small(self) -> bool:
        """Classifier-sized layers (<= 16 outputs) run on the CUDA-core kernels of csrc/ip_small.cu: a 128-row
        tensor-core tile would be mostly padding and three GEMM launches cost more than the arithmetic.  Not when the
        forward GEMM itself has to wait for the parameter server's version flags (non-rotated fused pull)."""
        # measured: 0.394 -> 0.387 ms per MNIST step (batch 1000); no gain at batch 64-128 (LeNet, CIFAR-10), where a
        # single 128-row tile already covers the batch
        return (getattr(self, "_ips_ws", None) is not None and self.vdim * (self.hdim | 1) * 4 <= 96 * 1024
                and self.pull_wait is None and os.environ.get("SINGA_B200_IP_SMALL", "1") != "0")

    def compute_feature(self, training):
        x = self._x2d()
        if self.native:
            K = _K()
            w16 = self._w16()
            if self._small():
                K.ip_small_fwd(x, w16, self.data, self.batchsize, self.vdim, self.hdim,
                               bias=self._bias_operand() if self.bias_term else None, act=self.fused_act)
            else:
                K.gemm(x, w16, self.batchsize, self.hdim, self.vdim, a_mn=False, b_mn=True, **_out(self.data),
                       bias=self._bias_operand() if self.bias_term else None, act=self.fused_act,
                       wait=self.pull_wait if training else None)
            self._x_saved = x
        else:
            y = R.ip_fwd(x, self.weight.data, self.bias.data if self.bias_term else None)
            if self.fused_act == ACT_RELU:
                y = R.relu_fwd(y)
            elif self.fused_act == ACT_STANH:
                y = R.stanh_fwd(y)
            self.data.copy_(y)
            self._x_saved = x

    split_backward = True

    def compute_gradient(self):
        self.compute_param_gradient()
        self.compute_data_gradient()

    def compute_param_gradient(self):
        """db = sum_rows(dY), dW = X^T dY (independent of the data gradient: may run on a side stream)."""
        dy, x = self.grad, self._x_saved
        if self.native:
            K = _K()
            accumulate = getattr(self, "_accumulate_wgrad", False)
            if self._small():     # weight and bias gradient in one launch
                K.ip_small_wgrad(x, dy, self.weight.grad, self.batchsize, self.vdim, self.hdim, self._ips_ws,
                                 db=self.bias.grad if self.bias_term else None, accumulate=accumulate)
                return
            # the (tiny) bias reduction goes first: issued behind the GEMM it would run while the previous bucket's
            # push kernel occupies the SMs, and the push of THIS layer waits for both
            if self.bias_term:
                K.colsum(dy, self.batchsize, self.hdim, dy.stride(0), self.bias.grad, accumulate=accumulate)
            # dW[vdim][hdim] = X^T dY : A = X stored [K=batch][M=vdim], B = dY stored [K=batch][N=hdim]
            K.gemm(x, dy, self.vdim, self.hdim, self.batchsize, a_mn=True, b_mn=True,
                   out32=self.weight.grad, accumulate=accumulate)
        else:
            _, dw, db = R.ip_bwd(x, self.weight.data, dy.float(), False)
            if getattr(self, "_accumulate_wgrad", False):
                self.weight.grad.add_(dw)
                if self.bias_term:
                    self.bias.grad.add_(db)
            else:
                self.weight.grad.copy_(dw)
                if self.bias_term:
                    self.bias.grad.copy_(db)

    def compute_data_gradient(self):
        """dX = dY W^T (times act'(x) when the source activation's backward is fused here)."""
        if not self._need_src_grad():
            return
        dy = self.grad
        dst, aux = self._src_grad()
        if self.native:
            K = _K()
            w16 = self._w16()
            if self.src4d:
                if self._small():
                    K.ip_small_dgrad(dy, w16, self.dx_flat, self.batchsize, self.vdim, self.hdim)
                else:
                    K.gemm(dy, w16, self.batchsize, self.vdim, self.hdim, **_out(self.dx_flat))
                n, c, h, w = self.srclayers[0].shape(self)
                if self.vdim % self.ldq == 0:
                    K.transpose_planes(self.dx_flat, dst, n, c, h * w)   # (c,h,w) -> NHWC
                else:
                    dst.copy_(self.dx_flat.reshape(dst.shape))
                if aux is not None:
                    K.act_bwd(dst, aux, dst, self.src_grad_act)
            else:
                d2 = self._flat2d(dst)
                a2 = self._flat2d(aux) if aux is not None else None
                if self._small():
                    K.ip_small_dgrad(dy, w16, d2, self.batchsize, self.vdim, self.hdim,
                                     grad_act=self.src_grad_act if a2 is not None else ACT_NONE, aux=a2)
                else:
                    K.gemm(dy, w16, self.batchsize, self.vdim, self.hdim, grad_act=self.src_grad_act, aux=a2,
                           **_out(d2))
        else:
            with R._mm_ctx(dy):
                dx = R._f(dy.float() @ self.weight.data.t())
            if aux is not None:
                a = aux.reshape(self.batchsize, self.vdim).float()
                dx = R.relu_bwd(dx, a) if self.src_grad_act == ACT_RELU else R.stanh_bwd(dx, a)
            dst.copy_(dx.view(dst.shape))


# ================================================================================================
# Convolution
# ================================================================================================
class ConvolutionLayer(NeuronLayer):
    """weight {F, C*k*k}, bias {F}; out = (H + 2p - k)/s + 1  (src/worker/layer.cc:18-123).

    Native path (``ops/conv.py``): implicit GEMM — ONE launch over all images per direction, the patch tiles built
    by TMA im2col loads inside the tcgen05 kernel, bias(+ReLU/tanh) in the forward epilogue, act' of the layer below
    in the data-gradient epilogue, channels-last output written directly."""
    type_id = "kConvolution"

    def connection_type(self, k):
        return S.kOneToAll

    def setup(self):
        assert len(self.srclayers) == 1
        cp = self.proto.convolution_param
        self.kernel, self.pad, self.stride = int(cp.kernel), int(cp.pad), int(cp.stride)
        assert self.kernel > 0, "Filter size cannot be zero."
        self.num_filters = int(cp.num_filters)
        if self.forced_shape is not None:
            self.num_filters = self.forced_shape[1]
            cp.num_filters = self.num_filters
        self.bias_term = cp.bias_term
        ss = self.srclayers[0].shape(self)
        assert len(ss) > 2, f"{self.name}: convolution needs an image source, got shape {ss}"
        self.width, self.height = ss[-1], ss[-2]
        self.channels = ss[-3] if len(ss) > 3 else 1
        self.batchsize = ss[0]
        self.conv_height = (self.height + 2 * self.pad - self.kernel) // self.stride + 1
        self.conv_width = (self.width + 2 * self.pad - self.kernel) // self.stride + 1
        self.col_height = self.channels * self.kernel * self.kernel
        self.col_width = self.conv_height * self.conv_width
        self._alloc([self.batchsize, self.num_filters, self.conv_height, self.conv_width])
        if not hasattr(self, "weight"):
            protos = list(self.proto.param)
            while len(protos) < 2:
                protos.append(S.ParamProto())
            if not protos[0].name:
                protos[0].name = "weight"
            if not protos[1].name:
                protos[1].name = "bias"
            self.weight = param_factory.create("Param")
            self.bias = param_factory.create("Param")
            self.weight.setup(protos[0], [self.num_filters, self.col_height], self.col_height)
            self.bias.setup(protos[1], [self.num_filters], 0)
        if self.native:
            self._setup_native()

    def _setup_native(self):
        from ..ops.conv import make_plan
        src = self.srclayers[0]
        need_dgrad = not (src.is_datalayer or src.is_parserlayer)
        self.P = self.batchsize * self.col_width
        self.plan = make_plan(self.batchsize, self.channels, self.height, self.width, self.num_filters, self.kernel,
                              self.stride, self.pad, ctx().act_dtype, ctx().device, need_dgrad=need_dgrad)
        self.weight.data16 = self.plan.wf      # the forward GEMM's filter operand ([F][(ky,kx,c)], zero padded)

        def pack(w=self.weight, plan=self.plan):
            # from the fp32 master copy, or — on a rank that is not this parameter's server — from the mirror the
            # parameter server broadcasts into
            plan.pack(w.mirror16 if getattr(w, "pack_from_mirror", False) and w.mirror16 is not None else w.data)
        self.weight.pack16 = pack

    def get_params(self):
        return [self.weight, self.bias] if self.bias_term else [self.weight]

    def _w16(self):
        w = self.weight if self.weight.is_owner else self.weight.owner
        return w.data16

    def _src_nhwc(self):
        src = self.srclayers[0].data_for(self)
        if src.dim() == 3:
            return src   # {N,H,W}: C == 1, contiguous == channels-last
        return src

    def compute_feature(self, training):
        src = self._src_nhwc()
        if self.native:
            from ..ops.conv import _mat
            y2d = _mat(self.data, self.P, self.num_filters, self.num_filters)    # view of the NHWC storage
            self.plan.forward(src, y2d, bias=self._bias_operand() if self.bias_term else None, act=self.fused_act)
        else:
            x = src.float()
            if x.dim() == 3:
                x = x.unsqueeze(1)
            y = R.conv_fwd(x, self.weight.data, self.bias.data if self.bias_term else None, self.kernel,
                           self.stride, self.pad)
            if self.fused_act == ACT_RELU:
                y = R.relu_fwd(y)
            elif self.fused_act == ACT_STANH:
                y = R.stanh_fwd(y)
            self.data.copy_(y)
            self._x_saved = x

    split_backward = True

    def compute_gradient(self):
        self.compute_param_gradient()
        self.compute_data_gradient()

    def compute_param_gradient(self):
        if self.native:
            accumulate = getattr(self, "_accumulate_wgrad", False)
            if self.bias_term:
                from ..ops.conv import _mat
                F_ = self.num_filters
                _K().colsum(_mat(self.grad, self.P, F_, F_), self.P, F_, F_, self.bias.grad, accumulate=accumulate)
            self.plan.wgrad(self.grad, self.weight.grad, accumulate=accumulate)
        else:
            self._dx_ref, dw, db = R.conv_bwd(self._x_saved, self.weight.data, self.grad.float(), self.kernel,
                                              self.stride, self.pad, self._need_src_grad())
            if getattr(self, "_accumulate_wgrad", False):
                self.weight.grad.add_(dw)
                if self.bias_term:
                    self.bias.grad.add_(db)
            else:
                self.weight.grad.copy_(dw)
                if self.bias_term:
                    self.bias.grad.copy_(db)

    def compute_data_gradient(self):
        if not self._need_src_grad():
            return
        dst, aux = self._src_grad()
        if self.native:
            from ..ops.conv import _mat
            rows, C = self.batchsize * self.height * self.width, self.channels
            self.plan.dgrad(self.grad, _mat(dst, rows, C, C), grad_act=self.src_grad_act,
                            aux2d=_mat(aux, rows, C, C) if aux is not None else None)
        else:
            dx = self._dx_ref
            if aux is not None:
                a = aux.float().view(dx.shape)
                dx = R.relu_bwd(dx, a) if self.src_grad_act == ACT_RELU else R.stanh_bwd(dx, a)
            dst.copy_(dx.view(dst.shape))


# ================================================================================================
# Activations
# ================================================================================================
class ActivationLayer(NeuronLayer):
    act = ACT_NONE

    def __init__(self):
        super().__init__()
        self.fwd_fused = False    # computed by the producer's epilogue
        self.bwd_fused = False    # computed by the consumer's dgrad epilogue

    def setup(self):
        assert len(self.srclayers) == 1
        src = self.srclayers[0]
        self.shape_ = src.shape(self)
        self.data = alloc_act(self.shape_)
        self.grad = alloc_act(self.shape_)

    def alias_to_source(self):
        """Forward fused into the producer: this layer's output IS the producer's output buffer, and the
        gradient w.r.t. the pre-activation is computed in place in the producer's gradient buffer."""
        src = self.srclayers[0]
        self.data = src.data_for(self)
        self.grad = src.grad_for(self)
        self.fwd_fused = True

    def compute_feature(self, training):
        if self.fwd_fused:
            return
        x = self.srclayers[0].data_for(self)
        if self.native:
            _K().act_fwd(x, self.data, self.act)
        else:
            self.data.copy_(R.relu_fwd(x) if self.act == ACT_RELU else R.stanh_fwd(x))

    def compute_gradient(self):
        if self.bwd_fused:
            return
        dst = self.srclayers[0].grad_for(self)
        if dst is None:
            return
        if self.native:
            _K().act_bwd(self.grad, self.data, dst, self.act)   # in place when grad aliases dst
        else:
            g, y = self.grad.float(), self.data.float()
            dst.copy_(R.relu_bwd(g, y) if self.act == ACT_RELU else R.stanh_bwd(g, y))


class ReLULayer(ActivationLayer):
    """max(x, 0); ``negative_slope`` is accepted and, as in the reference, unused (layer.cc:544-569)."""
    type_id = "kReLU"
    act = ACT_RELU


class TanhLayer(ActivationLayer):
    """1.7159047 tanh(0.66666667 x) — the reference hard-codes the scales (layer.cc:675-701)."""
    type_id = "kTanh"
    act = ACT_STANH


class DropoutLayer(NeuronLayer):
    """Training: y = x * mask / pkeep with a counter-RNG mask regenerated in backward (no mask tensor);
    test: identity (the reference applies the mask at test time too — survey §2.8 item 5 — not kept)."""
    type_id = "kDropout"

    def setup(self):
        assert len(self.srclayers) == 1
        self.pkeep = 1.0 - self.proto.dropout_param.dropout_ratio
        self._alloc(self.srclayers[0].shape(self))
        self.seed_base = (zlib.crc32(self.name.encode()) & 0xFFFFFF) << 20
        self.step_seed = 0
        self._training = False

    def set_step(self, step: int) -> None:
        self.step_seed = self.seed_base + step

    def compute_feature(self, training):
        x = self.srclayers[0].data_for(self)
        self._training = training
        if not training:
            self.data.copy_(x)
            return
        if self.native:
            _K().dropout(x, self.data, self.pkeep, self.seed_base, step_state()["step"])
        else:
            # same seed mixing as csrc/elementwise.cu::dropout_kernel, so both backends draw the same mask
            mixed = self.seed_base ^ (((int(step_state()['host_step'][0]) + 1) * 0x9E3779B97F4A7C15) & (2**64 - 1))
            m = R.dropout_mask(mixed, x.numel(), self.pkeep, x.device).view(x.shape)
            self._mask = m
            self.data.copy_(x.float() * m)

    def compute_gradient(self):
        dst = self.srclayers[0].grad_for(self)
        if dst is None:
            return
        if not self._training:
            dst.copy_(self.grad)
        elif self.native:
            _K().dropout(self.grad, dst, self.pkeep, self.seed_base, step_state()["step_bwd"])
        else:
            dst.copy_(self.grad.float() * self._mask)


# ================================================================================================
# Pooling / LRN
# ================================================================================================
class PoolingLayer(NeuronLayer):
    """MAX / AVE over k x k windows, stride s, output size in ceil form with clipped windows; AVE divides by
    k^2 (src/worker/layer.cc:476-540).  Padding is honoured (the reference reads ``pad_`` uninitialised)."""
    type_id = "kPooling"

    def setup(self):
        assert len(self.srclayers) == 1
        pp = self.proto.pooling_param
        self.kernel, self.stride, self.pad = int(pp.kernel), int(pp.stride), int(pp.pad)
        assert self.pad < self.kernel
        self.is_max = pp.pool == S.enum_value("PoolingProto.PoolMethod", "MAX")
        ss = self.srclayers[0].shape(self)
        assert len(ss) > 2
        self.width, self.height = ss[-1], ss[-2]
        self.channels = ss[-3] if len(ss) > 3 else 1
        self.batchsize = ss[0]
        self.pooled_height = R.pool_out_size(self.height, self.kernel, self.stride, self.pad)
        self.pooled_width = R.pool_out_size(self.width, self.kernel, self.stride, self.pad)
        self._alloc([self.batchsize, self.channels, self.pooled_height, self.pooled_width])

    def compute_feature(self, training):
        x = self.srclayers[0].data_for(self)
        if self.native:
            _K().pool_fwd(x, self.data, self.batchsize, self.height, self.width, self.channels,
                          self.pooled_height, self.pooled_width, self.kernel, self.stride, self.pad, self.is_max)
        else:
            x4 = x.float().view(self.batchsize, self.channels, self.height, self.width)
            self.data.copy_(R.pool_fwd(x4, self.kernel, self.stride, self.pad, self.is_max))

    def compute_gradient(self):
        src = self.srclayers[0]
        dst = src.grad_for(self)
        if dst is None:
            return
        x = src.data_for(self)
        if self.native:
            _K().pool_bwd(x, self.data, self.grad, dst, self.batchsize, self.height, self.width, self.channels,
                          self.pooled_height, self.pooled_width, self.kernel, self.stride, self.pad, self.is_max)
        else:
            x4 = x.float().view(self.batchsize, self.channels, self.height, self.width)
            dst.copy_(R.pool_bwd(x4, self.data.float(), self.grad.float(), self.kernel, self.stride, self.pad,
                                 self.is_max).view(dst.shape))


class LRNLayer(NeuronLayer):
    """Across-channel local response normalisation (src/worker/layer.cc:331-378).  WITHIN_CHANNEL is declared in
    the schema but, as in the reference, not implemented."""
    type_id = "kLRN"

    def setup(self):
        assert len(self.srclayers) == 1
        lp = self.proto.lrn_param
        self.lsize = int(lp.local_size)
        assert self.lsize % 2 == 1, "LRN only supports odd values for Localvol"
        if lp.norm_region != S.enum_value("LRNProto.NormRegion", "ACROSS_CHANNELS"):
            raise NotImplementedError("LRN WITHIN_CHANNEL is not implemented (neither is it in the reference)")
        self.alpha, self.beta, self.knorm = lp.alpha, lp.beta, lp.knorm
        ss = self.srclayers[0].shape(self)
        assert len(ss) == 4
        self.batchsize, self.channels, self.height, self.width = ss
        self._alloc(ss)
        self.norm = alloc_act(ss, torch.float32)

    def compute_feature(self, training):
        x = self.srclayers[0].data_for(self)
        if self.native:
            _K().lrn_fwd(x, self.data, self.norm, self.batchsize * self.height * self.width, self.channels,
                         self.lsize, self.alpha, self.beta, self.knorm)
        else:
            y, nm = R.lrn_fwd(x.float(), self.lsize, self.alpha, self.beta, self.knorm)
            self.data.copy_(y)
            self.norm.copy_(nm)

    def compute_gradient(self):
        src = self.srclayers[0]
        dst = src.grad_for(self)
        if dst is None:
            return
        x = src.data_for(self)
        if self.native:
            _K().lrn_bwd(x, self.norm, self.grad, dst, self.batchsize * self.height * self.width,
                         self.channels, self.lsize, self.alpha, self.beta)
        else:
            dst.copy_(R.lrn_bwd(x.float(), self.norm, self.grad.float(), self.lsize, self.alpha, self.beta))


# ================================================================================================
# Loss
# ================================================================================================
class LossLayer(NeuronLayer):
    is_losslayer = True

    def __init__(self):
        super().__init__()
        self.metric: Optional[torch.Tensor] = None


class SoftmaxLossLayer(LossLayer):
    """Softmax + NLL + top-k accuracy; metric = scale * [mean loss, accuracy]; gradient (p - onehot) scale / N
    written straight into the logits layer's gradient buffer in the forward pass
    (src/worker/layer.cc:703-764)."""
    type_id = "kSoftmaxLoss"

    @property
    def partition_type(self):
        pt = self.proto.partition_type if self.proto.HasField("partition_type") else S.kNone
        return S.kNone if pt == S.kLayerPartition else pt   # include/worker/layer.h:216-221

    def connection_type(self, k):
        return S.kOneToAll

    def setup(self):
        assert len(self.srclayers) == 2, f"{self.name}: softmax loss needs (logits, label) sources"
        ss = self.srclayers[0].shape(self)
        self.batchsize = ss[0]
        self.dim = 1
        for s in ss[1:]:
            self.dim *= s
        self.topk = int(self.proto.softmaxloss_param.topk)
        self.scale = float(self.proto.softmaxloss_param.scale)
        self.shape_ = [self.batchsize, self.dim]
        self.data = alloc_act(self.shape_, torch.float32)   # probabilities
        self.grad = None
        self.metric = torch.zeros(2, dtype=torch.float32, device=ctx().device)

    def compute_feature(self, training):
        logits = self.srclayers[0].data_for(self)
        labels = self.srclayers[1].data_for(self)
        # ``metric`` ACCUMULATES over steps on the device (Performance divides by its step counter and zeroes
        # it at the end of a display window), so no host round trip or memset sits in the step.
        if self.native:
            g = self.srclayers[0].grad_for(self) if training else None
            l2 = logits if logits.dim() == 2 else logits.view(self.batchsize, self.dim)
            if labels.dtype != torch.int32:
                # labels that travelled through Slice / Bridge layers of a partitioned net arrive as activations
                if getattr(self, "_labels_i32", None) is None:
                    self._labels_i32 = torch.zeros(self.batchsize, dtype=torch.int32, device=labels.device)
                self._labels_i32.copy_(labels.reshape(-1))
                labels = self._labels_i32
            _K().softmax_loss(l2, l2.stride(0), labels, self.batchsize, self.dim, self.topk, self.scale,
                              self.data, g, g.stride(0) if g is not None else 0, self.metric)
        else:
            m, g = R.softmax_loss(logits.float().view(self.batchsize, self.dim), labels, self.topk, self.scale)
            self.metric.add_(m)
            self._g = g

    def compute_gradient(self):
        if self.native:
            return   # gradient already written by the fused forward kernel
        dst = self.srclayers[0].grad_for(self)
        if dst is not None:
            dst.copy_(self._g.view(dst.shape))
