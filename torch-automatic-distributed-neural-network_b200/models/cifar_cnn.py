"""The reference example's CNN (cifar_example/sgd-torchad_nn-cifar.lua:88-118), B200 edition.

Architecture and parameter layout are the reference's: conv5x5(3->64)+ReLU+maxpool3/3,
conv5x5(64->64)+ReLU+maxpool3/3, View(64), Dropout(0.5), Linear(64,100)+ReLU, Linear(100,10),
LogSoftMax, trained with CrossEntropyCriterion (a second, idempotent log-softmax — SURVEY Q9);
114,838 parameters: ``conv1 [64,75]+64``, ``conv2 [64,1600]+64``, ``fc1 [100,64]+100``,
``fc2 [10,100]+10`` in Torch7 layouts, fp32 master copies.

Execution on CUDA (``loss(x, target)``): bf16 activations, NHWC between layers,
convolutions as im2col + the tcgen05 GEMM with bias+ReLU in its epilogue, pooling with saved
argmax, and one fused kernel for everything after the second pooling (dropout, both FC layers,
log-softmax, NLL and their gradients).  ``reference_logp`` is the plain PyTorch fp32 statement
of the same network used as the numerics oracle and as the CPU path.
"""
from __future__ import annotations

import ctypes as C
import math

import torch
import torch.nn.functional as F

from .. import _native
from ..nn.protocol import T7Protocol

K1PAD = 80  # 75 -> multiple of 8 (TMA pitch)

_SHADOWS: dict = {}
_DP1: dict = {}


def _shadows(dev, wkey: int):
    """Operand buffers of the fused path, per (device, model — identified by the address of its conv2
    weight): packed bf16 filters, the Toeplitz matrix of the transposed conv2, and the packed fp32 conv2
    weight-gradient scratch (kept zeroed by its consumer)."""
    key = (str(dev), int(wkey))
    sh = _SHADOWS.get(key)
    if sh is None:
        if len(_SHADOWS) > 8:
            _SHADOWS.clear()
        sh = {"w1p": torch.empty((64, 80), device=dev, dtype=torch.bfloat16),
              "w1x": torch.empty((128, 128), device=dev, dtype=torch.bfloat16),
              "w2p": torch.empty((64, 1600), device=dev, dtype=torch.bfloat16),
              "Tt": torch.empty((576, 4032), device=dev, dtype=torch.bfloat16),
              "dwp": torch.zeros((64, 1600), device=dev, dtype=torch.float32)}
        _SHADOWS[key] = sh
    return sh


def _toeplitz_kblock_mask(tile_n: int = 256):
    """Which 64-deep k-blocks (= window position pos of conv2's 3x3 pooled outputs) of the Toeplitz matrix
    ``Tt[pos*64 + o, (fy*9 + fx)*64 + c]`` are non-zero for each ``tile_n``-wide column tile: tap (kh, kw) =
    (fy - pos // 3, fx - pos % 3) must lie inside the 5x5 filter.  About half of the blocks are structural zeros."""
    px_per_tile = tile_n // 64
    masks = []
    for t in range((63 + px_per_tile - 1) // px_per_tile):
        m = 0
        for p in range(t * px_per_tile, min(63, (t + 1) * px_per_tile)):
            fy, fx = divmod(p, 9)
            for pos in range(9):
                if 0 <= fy - pos // 3 < 5 and 0 <= fx - pos % 3 < 5:
                    m |= 1 << pos
        masks.append(m)
    return masks


_TT_MASK = _toeplitz_kblock_mask(256)


def _dp1_buffer(dev, B):
    """[B,9,9,64] gradient buffer whose rows y >= 7 are zero and stay zero (the dgrad GEMM writes y < 7 only)."""
    key = (str(dev), int(B))
    t = _DP1.get(key)
    if t is None:
        if len(_DP1) > 4:
            _DP1.clear()
        t = torch.zeros((B, 9, 9, 64), device=dev, dtype=torch.bfloat16)
        _DP1[key] = t
    return t


def is_packed_input(x) -> bool:
    """NHWC4 bf16 [B,32,32,4] (channel 3 == 1.0): what ``CifarCNN.pack_input`` returns."""
    return x.dim() == 4 and tuple(x.shape[1:]) == (32, 32, 4) and x.dtype == torch.bfloat16


def _cnn_forward(x, target, w1, b1, w2, b2, fw1, fb1, fw2, fb2, p_drop, seed, fused_conv1, conv2_mode, seed_dev=None,
                 sliding=True):
    """Launch the forward kernels; returns (loss, logp, saved-for-backward dict)."""
    from ..ops import cnn as K
    B = x.shape[0]
    dev = x.device
    col1 = a1 = col2 = a2 = sh = None
    fused2 = conv2_mode == "dense" and fused_conv1
    sliding = bool(sliding) and fused2
    packed = is_packed_input(x)
    if packed and not sliding:
        raise ValueError("a packed (NHWC4) input needs the fused path with the sliding-operand conv1")
    if sliding:
        # conv1 reads the image itself as its tensor-core operand (csrc/conv1_sliding.cu): NHWC4 bf16, channel 3 == 1.0
        xb = x.contiguous() if packed else K.nhwc4_pack(x.contiguous())
        sh = _shadows(dev, w2.data_ptr())
        K.cnn_pack_shadows_sliding(w1, b1, w2, sh["w1x"], sh["w2p"], sh["Tt"])
        w2p = sh["w2p"]
        p1, idx1 = K.conv1s_pool_forward(xb, sh["w1x"])
        p2, idx2 = K.conv2_pool_forward(p1, w2p, b2)
    else:
        xb = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        xb = xb.contiguous()
    if sliding:
        pass
    elif fused2:
        # one launch refreshes every bf16 operand copy of the fp32 master weights (they may have been changed
        # by the optimiser, a synchronisation or a load since the last step)
        sh = _shadows(dev, w2.data_ptr())
        K.cnn_pack_shadows(w1, w2, sh["w1p"], sh["w2p"], sh["Tt"])
        w2p = sh["w2p"]
        p1, idx1 = K.conv1_pool_forward(xb, sh["w1p"], b1)
        # conv2 + bias + ReLU + pool2 as ONE implicit-GEMM kernel: 4-D TMA boxes of p1 are the A tiles
        p2, idx2 = K.conv2_pool_forward(p1, w2p, b2)
    else:
        w2p = K.conv_weight_pack(w2, 64, 5, 5, 1600)
        if fused_conv1:
            # implicit-GEMM conv + bias + ReLU + max-pool in one tcgen05 kernel: no col1 / a1 in HBM
            p1, idx1 = K.conv1_pool_forward(xb, K.conv1_weight_pack(w1), b1)
        else:
            w1p = K.conv_weight_pack(w1, 3, 5, 5, K1PAD)
            col1 = K.im2col(xb, 5, 5, K1PAD, nchw=True)                       # [B*784, 80]
            a1 = K.gemm_bf16(col1, w1p, bias=b1, act="relu", tile_n=64)       # [B*784, 64] == NHWC [B,28,28,64]
            p1, idx1 = K.maxpool_fwd(a1.view(B, 28, 28, 64), 3)               # [B,9,9,64]
        if conv2_mode == "legacy":
            col2 = K.im2col(p1, 5, 5, 1600, nchw=False)                       # [B*25, 1600]
            a2 = K.gemm_bf16(col2, w2p, bias=b2, act="relu", tile_n=64)       # [B*25, 64] == [B,5,5,64]
            p2, idx2 = K.maxpool_fwd(a2.view(B, 5, 5, 64), 3)                 # [B,1,1,64]
        else:
            # rows ordered with the 3x3 pixels that pool2 reads first; "dense" still evaluates the
            # convolution at all 25 pixels (as the reference does), "pooled" only at those 9
            col2 = K.im2col_winfirst(p1, 5, 5, 3, only_window=(conv2_mode == "pooled"))
            a2 = K.gemm_bf16(col2, w2p, bias=b2, act="relu", tile_n=64)
            p2, idx2 = K.maxpool_fwd(a2[:B * 9].view(B, 3, 3, 64), 3)         # [B,1,1,64]
    hidden = torch.empty((B, 100), device=dev, dtype=torch.float32)
    logp = torch.empty((B, 10), device=dev, dtype=torch.float32)
    loss = torch.empty((), device=dev, dtype=torch.float32)
    tgt = target.to(device=dev, dtype=torch.int64).contiguous()
    seed = int(seed) & 0xFFFFFFFF
    h = K.HeadArgs()
    h.p2, h.w1, h.b1, h.w2, h.b2 = p2.data_ptr(), fw1.data_ptr(), fb1.data_ptr(), fw2.data_ptr(), fb2.data_ptr()
    h.target = tgt.data_ptr()
    h.B, h.D, h.H, h.NC = B, 64, 100, 10
    h.dropout_p, h.seed = float(p_drop), seed
    if seed_dev is not None:
        h.seed_dev = seed_dev.data_ptr()
    h.hidden, h.logp, h.loss = hidden.data_ptr(), logp.data_ptr(), loss.data_ptr()
    ws = K.head_workspace(dev)
    _native.check(_native.lib().tadnn_head_forward(C.byref(h), C.c_void_p(ws.data_ptr()),
                                                   C.c_void_p(_native.stream_ptr())), "tadnn_head_forward")
    saved = dict(xb=xb, col1=col1, a1=a1, p1=p1, idx1=idx1, col2=col2, a2=a2, p2=p2, idx2=idx2, hidden=hidden,
                 logp=logp, tgt=tgt, w2p=w2p, fw1=fw1, fb1=fb1, fw2=fw2, fb2=fb2, p_drop=float(p_drop), seed=seed, seed_dev=seed_dev,
                 B=B, fused_conv1=bool(fused_conv1), conv2_mode=conv2_mode, fused2=fused2, sh=sh, sliding=sliding)
    return loss, logp, saved


def _cnn_backward(sv, gw1, gb1, gw2, gb2, gfw1, gfb1, gfw2, gfb2, gloss=None, tail_ready=None):
    """Launch the backward kernels; every gradient is ACCUMULATED (+=) into the given fp32 buffers
    (Torch7 layouts).  ``gloss``: optional device scalar dLoss."""
    from ..ops import cnn as K
    B, dev = sv["B"], sv["p2"].device
    p1, idx1, p2, idx2, col2, a2, w2p = sv["p1"], sv["idx1"], sv["p2"], sv["idx2"], sv["col2"], sv["a2"], sv["w2p"]
    fused2 = sv.get("fused2", False)
    h = K.HeadArgs()
    h.p2, h.w1, h.b1, h.w2, h.b2 = p2.data_ptr(), sv["fw1"].data_ptr(), sv["fb1"].data_ptr(), sv["fw2"].data_ptr(), sv["fb2"].data_ptr()
    h.target = sv["tgt"].data_ptr()
    h.B, h.D, h.H, h.NC = B, 64, 100, 10
    h.dropout_p, h.seed = sv["p_drop"], sv["seed"]
    if sv.get("seed_dev") is not None:
        h.seed_dev = sv["seed_dev"].data_ptr()
    h.hidden, h.logp = sv["hidden"].data_ptr(), sv["logp"].data_ptr()
    h.gw1, h.gb1, h.gw2, h.gb2 = gfw1.data_ptr(), gfb1.data_ptr(), gfw2.data_ptr(), gfb2.data_ptr()
    h.grad_scale = 1.0
    if gloss is not None:
        gl = gloss.to(device=dev, dtype=torch.float32).contiguous()
        h.grad_scale_dev = gl.data_ptr()  # upstream dLoss stays on the device: no host sync
    if fused2:
        # head backward also routes the gradient through pool2 + ReLU (-> da2w, rows (n, window pixel)) and
        # takes conv2's bias gradient; conv2's weight gradient reads p1 through TMA boxes (no im2col);
        # the transposed convolution is one GEMM against the Toeplitz matrix, written straight into the
        # full-size dp1 (its never-written rows stay zero); conv1's backward also takes conv1's bias gradient.
        sh = sv["sh"]
        da2w = torch.empty((B * 9, 64), device=dev, dtype=torch.bfloat16)
        h.pool_idx, h.dpre, h.gbias_pre = idx2.data_ptr(), da2w.data_ptr(), gb2.data_ptr()
        _native.check(_native.lib().tadnn_head_backward(C.byref(h), C.c_void_p(_native.stream_ptr())), "tadnn_head_backward")
        K.conv2_wgrad(p1, da2w, sh["dwp"])
        K.conv2_wgrad_unpack(sh["dwp"], gw2)
        dp1 = _dp1_buffer(dev, B)
        K.gemm_bf16(da2w.view(B, 576), sh["Tt"], b_mn_major=True, out=dp1.view(B, 81 * 64)[:, :7 * 9 * 64], tile_n=256,
                    kblock_mask=_TT_MASK)
        if tail_ready is not None:
            tail_ready()  # every gradient except conv1's is complete (and nothing below reads those parameters)
        if sv.get("sliding"):
            K.conv1s_pool_backward(sv["xb"], dp1, idx1, gw1, gb1, rows=7)   # dp1 rows 7, 8 are never written (zero)
        else:
            K.conv1_pool_backward_bias(sv["xb"], dp1, p1, idx1, gw1, gb1)
        return
    dp2 = torch.empty((B, 1, 1, 64), device=dev, dtype=torch.bfloat16)
    h.dp2 = dp2.data_ptr()
    _native.check(_native.lib().tadnn_head_backward(C.byref(h), C.c_void_p(_native.stream_ptr())), "tadnn_head_backward")
    K.colsum_relu_mask(dp2, p2, gb2)
    if sv["conv2_mode"] == "legacy":
        # pool/relu backward into a2's storage, wgrad (split-K), dgrad into col2's storage, col2im
        da2 = K.maxpool_relu_bwd(dp2, p2, idx2, 5, 5, 3, out=a2.view(B, 5, 5, 64)).view(B * 25, 64)
        dw2p = K.gemm_bf16(da2, col2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                           split_k=K.auto_split_k(64, 1600, B * 25))
        K.conv_wgrad_unpack(dw2p, gw2, 64, 5, 5)
        dcol2 = K.gemm_bf16(da2, w2p, b_mn_major=True, out=col2)
        dp1 = K.col2im(dcol2, B, 64, 9, 9, 5, 5)
        K.colsum_relu_mask(dp1, p1, gb1)
    else:
        # d(conv2 out) is exactly zero outside the 3x3 pooled window: wgrad over those 9 pixels'
        # patches, and the transposed convolution as ONE dense GEMM against the Toeplitz-expanded
        # weights ([B, 9*64] x [9*64, 7*7*64]) instead of a 25-tap dgrad GEMM + col2im.  The result
        # is the 7x7 top-left corner of dp1 (the rest is zero) and is consumed in that compact form.
        da2w = K.maxpool_relu_bwd(dp2, p2, idx2, 3, 3, 3).view(B * 9, 64)
        dw2p = K.gemm_bf16(da2w, col2[:B * 9], a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                           split_k=K.auto_split_k(64, 1600, B * 9))
        K.conv_wgrad_unpack(dw2p, gw2, 64, 5, 5)
        T = K.conv_toeplitz_pack(w2p, 64, 64, 5, 5, 3)                    # [7*7*64, 9*64]
        dp1 = K.gemm_bf16(da2w.view(B, 576), T).view(B, 7, 7, 64)        # compact corner of [B,9,9,64]
        K.colsum_relu_mask_corner(dp1, p1, gb1)
        dp1 = K.expand_corner(dp1, 9, 9)  # (the fused conv1 backward also accepts the compact corner; full is faster)
    # conv1 (no dgrad: the image needs no gradient)
    if sv["fused_conv1"]:
        K.conv1_pool_backward(sv["xb"], dp1, p1, idx1, gw1)
    else:
        a1, col1 = sv["a1"], sv["col1"]
        da1 = K.maxpool_relu_bwd(dp1, p1, idx1, 28, 28, 3, out=a1.view(B, 28, 28, 64)).view(B * 784, 64)
        dw1p = K.gemm_bf16(da1, col1, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                           split_k=K.auto_split_k(64, K1PAD, B * 784, 128), tile_n=128)
        K.conv_wgrad_unpack(dw1p, gw1, 3, 5, 5)


class _CnnLoss(torch.autograd.Function):
    """Autograd wrapper (``loss.backward()`` idiom).  ``CifarCNN.forward_backward`` is the same pair
    of launch sequences without autograd, accumulating straight into ``param.grad``."""

    @staticmethod
    def forward(ctx, x, target, w1, b1, w2, b2, fw1, fb1, fw2, fb2, p_drop, seed, fused_conv1, conv2_mode, sliding=True):
        loss, logp, sv = _cnn_forward(x, target, w1, b1, w2, b2, fw1, fb1, fw2, fb2, p_drop, seed, fused_conv1, conv2_mode,
                                      sliding=sliding)
        ctx.sv = sv
        ctx.mark_non_differentiable(logp)
        return loss, logp

    @staticmethod
    def backward(ctx, gloss, _glogp):
        sv = ctx.sv
        dev = sv["p2"].device
        sizes = [(64, 75), (64,), (64, 1600), (64,), (100, 64), (100,), (10, 100), (10,)]
        flat = torch.zeros(sum(math.prod(s) for s in sizes), device=dev, dtype=torch.float32)  # one fill
        grads, off = [], 0
        for sz in sizes:
            n = math.prod(sz)
            grads.append(flat[off:off + n].view(sz))
            off += n
        _cnn_backward(sv, *grads, gloss=gloss)
        ctx.sv = None
        return (None, None, *grads, None, None, None, None, None)


class CifarCNN(torch.nn.Module, T7Protocol):
    # uint8 pixels -> roughly unit-variance bf16: what BatchTrainer.fit applies on the device when the
    # dataset is shipped as bytes (the reference's own preprocessing is utils/preprocess.py)
    input_affine = ([4.0 / 255.0] * 3, [-2.0] * 3)

    def __init__(self, dropout: float = 0.5, seed: int = 1, fused_conv1: bool = True, conv2_mode: str = "dense",
                 conv1_kernel: str = "sliding"):
        super().__init__()
        assert conv2_mode in ("dense", "pooled", "legacy")
        assert conv1_kernel in ("sliding", "rows")
        self.dropout = dropout
        self.fused_conv1 = fused_conv1  # False: im2col + GEMM + separate pooling kernels (generic path)
        # fused conv1 kernels: "sliding" = the staged NHWC4 image is the tensor-core operand (csrc/conv1_sliding.cu; fused
        # "dense" path only), "rows" = one im2col row per pixel built in shared memory (csrc/conv1_fused.cu)
        self.conv1_kernel = conv1_kernel
        # conv2 forward: "dense" evaluates all 5x5 output pixels like the reference; "pooled" only the
        # 3x3 that pool2 reads (identical loss and gradients); both use the sparse-gradient backward.
        # "legacy": raster-order im2col, 25-tap dgrad GEMM + col2im (first implementation, kept as oracle).
        self.conv2_mode = conv2_mode
        self._step = 0
        self._seed = seed
        P = torch.nn.Parameter
        self.conv1_weight, self.conv1_bias = P(torch.empty(64, 75)), P(torch.empty(64))
        self.conv2_weight, self.conv2_bias = P(torch.empty(64, 1600)), P(torch.empty(64))
        self.fc1_weight, self.fc1_bias = P(torch.empty(100, 64)), P(torch.empty(100))
        self.fc2_weight, self.fc2_bias = P(torch.empty(10, 100)), P(torch.empty(10))
        self.reset()

    def reset(self):
        """Torch7 resets: U(+-1/sqrt(fan_in)) for weight and bias of every layer."""
        with torch.no_grad():
            for w, b, fan in ((self.conv1_weight, self.conv1_bias, 75), (self.conv2_weight, self.conv2_bias, 1600),
                              (self.fc1_weight, self.fc1_bias, 64), (self.fc2_weight, self.fc2_bias, 100)):
                s = 1.0 / math.sqrt(fan)
                w.uniform_(-s, s)
                b.uniform_(-s, s)
        return self

    # ---- plain PyTorch statement of the network (oracle / CPU path) -----------------------
    def reference_logp(self, x: torch.Tensor, dropout_scale: torch.Tensor | None = None,
                       emulate_bf16: bool = False) -> torch.Tensor:
        """``emulate_bf16`` rounds weights and post-ReLU activations to bf16 at the points where
        the fused path stores bf16, so that pooling picks the same arg-max in both."""
        def r(t):
            return t.to(torch.bfloat16).float() if emulate_bf16 else t
        x = x.float()
        # the fused conv1 carries its bias through the GEMM as a bf16 operand column
        y = F.conv2d(x, r(self.conv1_weight).view(64, 3, 5, 5), r(self.conv1_bias) if self.fused_conv1 else self.conv1_bias)
        # the sliding-operand conv1 pools the fp32 accumulators and rounds the winner (same values, since rounding is
        # monotonic, but a tie between two values that round to the same bf16 goes to the larger fp32 one, not to the first)
        y = r(F.max_pool2d(F.relu(y), 3, 3)) if self.packs_input else F.max_pool2d(r(F.relu(y)), 3, 3)
        y = F.conv2d(y, r(self.conv2_weight).view(64, 64, 5, 5), self.conv2_bias)
        y = F.max_pool2d(r(F.relu(y)), 3, 3).reshape(-1, 64)
        if dropout_scale is not None:
            y = y * dropout_scale
        elif self.training and self.dropout > 0:
            y = F.dropout(y, self.dropout, True)
        y = F.relu(F.linear(y, self.fc1_weight, self.fc1_bias))
        y = F.linear(y, self.fc2_weight, self.fc2_bias)
        return F.log_softmax(y, dim=-1)

    def _native_ok(self, x) -> bool:
        return x.is_cuda and x.dim() == 4 and (tuple(x.shape[1:]) == (3, 32, 32) or (self.packs_input and is_packed_input(x)))

    @property
    def packs_input(self) -> bool:
        """True when the fused path reads its input as NHWC4 bf16 (``pack_input``)."""
        return self.fused_conv1 and self.conv2_mode == "dense" and self.conv1_kernel == "sliding"

    def pack_input(self, x: torch.Tensor, affine=None, out=None) -> torch.Tensor:
        """[B,3,32,32] uint8 / bf16 / fp32 on the device -> the fused path's own input layout, NHWC4 bf16 with channel
        3 == 1.0 (the bias rides on it), ``y = x * scale[c] + shift[c]`` for ``affine = (scale, shift)``.  Every entry point
        also takes the plain NCHW tensor and packs it itself (one more launch per step); a data pipeline that packs once
        (``BatchTrainer.fit`` does, in the kernel that widens the uint8 bytes) saves that launch."""
        from ..ops import cnn as K
        sc, sh = affine if affine is not None else (None, None)
        return K.nhwc4_pack(x.contiguous(), sc, sh, out=out)

    def loss_and_logp(self, x: torch.Tensor, target: torch.Tensor):
        """Mean cross-entropy (differentiable) and log-probabilities."""
        if x.dim() == 3:
            x, target = x.unsqueeze(0), torch.as_tensor(target).reshape(1)
        if self._native_ok(x):
            self._step += 1
            p = self.dropout if self.training else 0.0
            seed = (self._seed * 2654435761 + self._step * 40503) & 0xFFFFFFFF
            return _CnnLoss.apply(x, target.to(x.device), self.conv1_weight, self.conv1_bias, self.conv2_weight,
                                  self.conv2_bias, self.fc1_weight, self.fc1_bias, self.fc2_weight, self.fc2_bias, p, seed,
                                  self.fused_conv1, self.conv2_mode, self.conv1_kernel == "sliding")
        logp = self.reference_logp(x)
        return F.nll_loss(logp, target.long().to(logp.device)), logp.detach()

    def loss(self, x, target):
        return self.loss_and_logp(x, target)[0]

    def _plist(self):
        return [self.conv1_weight, self.conv1_bias, self.conv2_weight, self.conv2_bias,
                self.fc1_weight, self.fc1_bias, self.fc2_weight, self.fc2_bias]

    def forward_backward(self, x: torch.Tensor, target: torch.Tensor, tail_ready=None) -> torch.Tensor:
        """Torch7-style explicit step (``model:forward`` + ``model:backward``): runs the fused
        forward and backward kernels without autograd and ACCUMULATES the gradients into
        ``param.grad`` (= gradWeight/gradBias).  Returns the loss tensor (device).  The caller runs
        the synchronisation hook (``model._tadnn_sync.after_backward(model)``), as the hooked
        ``Sequential:backward`` does (datamodule.lua:96-110).  ``tail_ready(first_param)``, if given, is called
        from inside the backward pass once the gradients of ``first_param`` and of every parameter after it
        are final (here: everything but conv1), so that their synchronisation can overlap conv1's backward."""
        if not self._native_ok(x):
            loss = self.loss(x, target)
            loss.backward()
            return loss.detach()
        ps = self._plist()
        for p in ps:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        self._step += 1
        pd = self.dropout if self.training else 0.0
        seed = (self._seed * 2654435761 + self._step * 40503) & 0xFFFFFFFF
        if self.packs_input and all(p.dtype == torch.float32 and p.grad.dtype == torch.float32 for p in ps):
            return self._native_step(x, target, ps, pd, seed, tail_ready)
        with torch.no_grad():
            loss, _logp, sv = _cnn_forward(x, target, *[p.data for p in ps], pd, seed, self.fused_conv1, self.conv2_mode,
                                           seed_dev=getattr(self, "_seed_dev", None), sliding=self.conv1_kernel == "sliding")
            cb = (lambda: tail_ready(self.conv2_weight)) if (tail_ready is not None and sv.get("fused2")) else None
            _cnn_backward(sv, *[p.grad for p in ps], tail_ready=cb)
        return loss

    def _native_step(self, x, target, ps, pd, seed, tail_ready):
        """The fused step through ``tadnn_cnn_step`` (csrc/cnn_step.cu): one or two native calls launch all ten kernels;
        activations live in a workspace kept per (device, batch size)."""
        from ..ops import cnn as K
        dev, B = x.device, x.shape[0]
        xq = x.contiguous() if is_packed_input(x) else self.pack_input(x)
        ws = self.__dict__.get("_step_ws")
        if ws is None or ws["key"] != (dev, B, self.conv2_weight.data_ptr()):
            sh = _shadows(dev, self.conv2_weight.data_ptr())
            bf, u8, f32 = torch.bfloat16, torch.uint8, torch.float32
            t = {"p1": torch.empty((B, 9, 9, 64), device=dev, dtype=bf), "idx1": torch.empty((B, 9, 9, 64), device=dev, dtype=u8),
                 "p2": torch.empty((B, 64), device=dev, dtype=bf), "idx2": torch.empty((B, 64), device=dev, dtype=u8),
                 "hidden": torch.empty((B, 100), device=dev, dtype=f32), "logp": torch.empty((B, 10), device=dev, dtype=f32),
                 "da2w": torch.empty((B * 9, 64), device=dev, dtype=bf), "dp1": _dp1_buffer(dev, B), "head_ws": K.head_workspace(dev),
                 "w1x": sh["w1x"], "w2p": sh["w2p"], "Tt": sh["Tt"], "dwp": sh["dwp"]}
            a = K.CnnStepArgs()
            for n, v in t.items():
                setattr(a, n, v.data_ptr())
            a.B = B
            a.grad_scale = 1.0
            a.tt_mask = (C.c_uint * len(_TT_MASK))(*_TT_MASK)
            a.n_tt_mask = len(_TT_MASK)
            ws = {"key": (dev, B, self.conv2_weight.data_ptr()), "t": t, "a": a}
            self.__dict__["_step_ws"] = ws
        a = ws["a"]
        tgt = target if (target.dtype == torch.int64 and target.device == dev and target.is_contiguous()) \
            else target.to(device=dev, dtype=torch.int64).contiguous()
        loss = torch.empty((), device=dev, dtype=torch.float32)
        a.xq, a.target, a.loss = xq.data_ptr(), tgt.data_ptr(), loss.data_ptr()
        a.w1, a.b1, a.w2, a.b2, a.fw1, a.fb1, a.fw2, a.fb2 = [p.data_ptr() for p in ps]
        a.gw1, a.gb1, a.gw2, a.gb2, a.gfw1, a.gfb1, a.gfw2, a.gfb2 = [p.grad.data_ptr() for p in ps]
        a.dropout_p, a.seed = float(pd), int(seed)
        sd = getattr(self, "_seed_dev", None)
        a.seed_dev = sd.data_ptr() if sd is not None else None
        a.forward_only = 0
        if tail_ready is None:
            a.phase = 3
            K.cnn_step(a)
        else:
            a.phase = 1
            K.cnn_step(a)
            tail_ready(self.conv2_weight)   # every gradient except conv1's is complete
            a.phase = 2
            K.cnn_step(a)
        ws["keep"] = (xq, tgt)   # inputs of the launches just queued: alive until the next step replaces them
        return loss

    # ---- the reference's own workload: one sample per step (datamodule.lua:134-158) ----------------
    def native_sample_step(self, x: torch.Tensor, target, lr: float, weight_decay: float = 0.0) -> torch.Tensor:
        """Forward + criterion + backward + in-place SGD step for ONE sample (or a small batch) as a replay of
        a captured CUDA graph: the ten kernels of the fused step plus the optimiser launch cost one graph
        launch instead of eleven launches and their Python glue, which is what bounds a B = 1 loop.
        The dropout mask of each replay comes from a counter in device memory that the graph itself
        increments.  Returns the (device) loss; never synchronises.  Used by ``nn.StochasticGradient``."""
        if x.dim() == 3:
            x = x.unsqueeze(0)
        tgt = torch.as_tensor(target, device=x.device).reshape(-1)
        key = (tuple(x.shape), float(lr), float(weight_decay), self.training, self.conv1_weight.data_ptr())
        ent = self.__dict__.get("_sample_graph")
        if ent is None or ent["key"] != key:
            ent = self._capture_sample_graph(x, tgt, lr, weight_decay, key)
            self.__dict__["_sample_graph"] = ent
        ent["x"].copy_(x, non_blocking=True)
        ent["y"].copy_(tgt, non_blocking=True)
        ent["graph"].replay()
        return ent["loss"]

    def _capture_sample_graph(self, x, tgt, lr, wd, key):
        from ..ops.cnn import sgd_step
        dev = x.device
        xs = torch.zeros(x.shape, device=dev, dtype=torch.bfloat16)
        ys = torch.zeros(tgt.shape, device=dev, dtype=torch.int64)
        self._seed_dev = torch.zeros(1, device=dev, dtype=torch.int32)
        ps = self._plist()
        for p in ps:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        flat = getattr(self, "_tadnn_flat", None)

        def body(lr_, wd_):
            loss = self.forward_backward(xs, ys)
            if flat is not None and all(dt == torch.float32 for dt in flat.groups):
                for g in flat.groups.values():
                    sgd_step(g.region[:g.half], g.region[g.half:], lr_, wd_, zero_grad=True)
            else:
                for p in ps:
                    sgd_step(p.data.view(-1), p.grad.view(-1), lr_, wd_, zero_grad=True)
            self._seed_dev.add_(1)
            return loss

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):   # warm-up outside capture: lazy one-time set-up (smem attributes, workspaces, buffers)
            xs.copy_(x)
            ys.copy_(tgt)
            for _ in range(2):
                body(0.0, 0.0)   # zero step size: the warm-up must leave the parameters as they are
            self._seed_dev.zero_()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            loss = body(lr, wd)
        return {"key": key, "graph": graph, "x": xs, "y": ys, "loss": loss}

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """Log-probabilities ``[B, 10]`` (what the reference model outputs after LogSoftMax)."""
        squeeze = x.dim() == 3
        if squeeze:
            x = x.unsqueeze(0)
        if self._native_ok(x) and not torch.is_grad_enabled():
            dummy = torch.zeros(x.shape[0], dtype=torch.int64, device=x.device)
            was = self.training
            self.training = False
            try:
                logp = self.loss_and_logp(x, dummy)[1]
            finally:
                self.training = was
        else:
            logp = self.reference_logp(x)
        return logp.squeeze(0) if squeeze else logp

    def to_sequential(self):
        """The same network as a ``tadnn.nn.Sequential`` of Torch7-named layers sharing these
        parameters (for ``.t7`` export and for the per-sample reference-config example)."""
        from .. import nn as tnn
        c1, c2 = tnn.SpatialConvolutionMM(3, 64, 5, 5), tnn.SpatialConvolutionMM(64, 64, 5, 5)
        l1, l2 = tnn.Linear(64, 100), tnn.Linear(100, 10)
        c1.weight, c1.bias, c2.weight, c2.bias = self.conv1_weight, self.conv1_bias, self.conv2_weight, self.conv2_bias
        l1.weight, l1.bias, l2.weight, l2.bias = self.fc1_weight, self.fc1_bias, self.fc2_weight, self.fc2_bias
        return (tnn.Sequential().add(c1).add(tnn.ReLU()).add(tnn.SpatialMaxPooling(3, 3, 3, 3))
                .add(c2).add(tnn.ReLU()).add(tnn.SpatialMaxPooling(3, 3, 3, 3)).add(tnn.View(64))
                .add(tnn.Dropout(self.dropout)).add(l1).add(tnn.ReLU()).add(l2).add(tnn.LogSoftMax()))


class CnnCriterion:
    """Pairs with ``CifarCNN`` in ``StochasticGradient``-style loops: ``crit(model_out, y)`` needs
    the model's *input*, so the fused path is exposed as ``model.loss(x, y)`` instead; this class
    only adapts the ``criterion(output, target)`` call for log-prob outputs."""

    def __call__(self, logp, target):
        return F.nll_loss(logp.float(), torch.as_tensor(target).long().reshape(-1).to(logp.device))
