"""Fused model-parallel MLP: the reference's ``MPInitialLinear -> MPTanh -> MPBaseLinear ...``
stack (README.md:97-101; nodemodule.lua:34-174) as B200 kernels.

Parameter layout is the reference's: rank r holds ``W_r [o, i/P]`` (input-dimension split) and a
full ``bias [o]``; the effective dense layer is ``W = cat_r W_r``, ``b = sum_r b_r``.

Dataflow (what the reference does with a cuBLAS SGEMM + blocking ``allreduceTensor`` per layer
on replicated activations) becomes, per layer::

    forward   partial_r = a_r · W_rᵀ + b_r           tcgen05 GEMM; its epilogue TMA-stores each
                                                      64-column slab straight into the OWNER
                                                      rank's staging slot over NVLink
              a'_r = tanh( sum_q slot_q )             reduce_slots kernel (barrier + local reduce)
    backward  g = all_gather(g_r)                     push all-gather (csrc/comm.cu)
              dW_r = gᵀ · a_r ,  db = colsum(g)       tcgen05 GEMM (MN-major operands, split-K)
              g'_r = (g · W_r) ⊙ (1 - a_r²)           tcgen05 GEMM, tanh' fused in the epilogue

Activations stay *sharded* between layers (``a'_r`` is exactly the slice the next layer's
``narrowInput`` would take, nodemodule.lua:16-21), so the forward moves (P-1)/P·B·o elements per
layer instead of the all-reduce's 2(P-1)/P·B·o, and the redundant full-width tanh of the
reference disappears.  The backward is the exact gradient (the reference's ``syncTanh`` gather).
Layers whose width is not a multiple of 64·P (e.g. the 10-way classifier) fall back to a plain
GEMM + one-shot all-reduce of the small output; a classifier-sized layer (o <= 16) runs on the
GEMV-style kernels of csrc/smalln.cu instead of a mostly empty MMA tile.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional

import torch
import torch.nn.functional as F

from .. import _native
from .. import mpi as _mpi
from . import symm as _symm
from ..ops import smalln as _smalln


def _cast_bf16(w: torch.Tensor) -> torch.Tensor:
    out = torch.empty(w.shape, device=w.device, dtype=torch.bfloat16)
    _native.check(_native.lib().tadnn_cast_f32_bf16(C.c_void_p(w.data_ptr()), C.c_void_p(out.data_ptr()),
                                                    C.c_longlong(w.numel()), C.c_void_p(_native.stream_ptr())),
                  "tadnn_cast_f32_bf16")
    return out


def _smalln_ok(o: int, a: torch.Tensor, which: str) -> bool:
    """The GEMV-style kernels (csrc/smalln.cu) take a classifier-sized layer when the activation rows are
    16-byte aligned and the measured crossover (ops/smalln.py) says they beat the tcgen05 tile."""
    prefer = getattr(_smalln, "prefer_" + which)
    return (prefer(a.shape[0], o, a.shape[1]) and a.stride(1) == 1 and a.stride(0) % 8 == 0
            and a.data_ptr() % 16 == 0)


class _Plan:
    """Per (batch size) symmetric buffers shared by all layers of one FusedMPMLP."""

    def __init__(self, ctx: _symm.SymmContext, B: int, sizes: List[int]):
        P = ctx.world
        self.B = B
        big = [o for o in sizes[1:] if o % (64 * P) == 0]
        maxo = max(big) if big else 0
        allo = max(sizes[1:])
        allo_pad = (allo + 7) // 8 * 8
        self.Bpad = (B + 31) // 32 * 32
        self.slots_off = ctx.alloc(2 * self.Bpad * maxo) if maxo else 0  # [P] slots of Bpad x o/P bf16 (slab-major)
        self.gfull_off = ctx.alloc(2 * B * allo_pad)                  # [B, o] bf16 (gathered grads / outputs)
        self.gfull2_off = ctx.alloc(2 * B * allo_pad)                 # second buffer: the all-gather of layer l-1's
        #                                                               gradient lands here while layer l's wgrad still reads the first
        self.small = ctx.empty((B * allo_pad,), torch.float32)        # small-layer partial sums (all-reduced in place)
        self._ctx, self._maxo = ctx, maxo
        self._part2 = None
        # fused all-gather -> GEMM backward (K13): three gathered-gradient buffers (layer l uses l % 3: a peer may
        # already push layer l-1's slices while this rank still reads layer l+1's) and, per buffer, one
        # arrival counter per (32-row block, source rank); counters only grow, `gtarget` is their running goal
        self.g3_off = None
        self.gflag_off = None
        self.gtarget = [0, 0, 0]
        self._B, self._allo_pad = B, allo_pad

    def gather_buffers(self):
        if self.g3_off is None:
            ctx, P = self._ctx, self._ctx.world
            self.g3_off = [ctx.alloc(2 * self._B * self._allo_pad) for _ in range(3)]
            nflag = (self._B + 31) // 32 * P
            self.gflag_off = [ctx.alloc(4 * nflag) for _ in range(3)]
            for off in self.gflag_off:
                ctx.tensor(off, (nflag,), torch.float32).zero_()
            torch.cuda.synchronize()
            if P > 1:
                import torch.distributed as dist
                dist.barrier()  # nobody signals into counters that are not zeroed yet
        return self.g3_off, self.gflag_off

    def part_off(self, parity: int) -> int:
        """NVLS mode: [B, o] bf16 partial-product buffers, double-buffered over layers (a peer may still be
        reading layer l's partials through the switch while this rank's layer l+1 GEMM already writes)."""
        if parity == 0:
            return self.slots_off
        if self._part2 is None:
            self._part2 = self._ctx.alloc(2 * self.Bpad * self._maxo)
        return self._part2


class _Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, mod: "FusedMPMLP", x: torch.Tensor, *params):
        from ..ops.gemm import gemm_bf16
        sc = mod._ctx()
        P, r = sc.world, sc.rank
        L = mod.num_layers
        B = x.shape[0]
        plan = mod._plan(B)
        lib = _native.lib()
        st = C.c_void_p(_native.stream_ptr())
        xb = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        w0 = mod.sizes[0] // P
        a = xb[:, r * w0:(r + 1) * w0]  # narrowInput: a strided view, consumed in place by TMA
        acts = [a]
        wbs = []
        out = None
        use_nvls = P > 1 and sc.has_multicast and mod.reduce_mode == "nvls"
        trace = getattr(mod, "_trace", None)  # tools/mp_layer_timeline.py: a list that collects (label, event)

        def mark(label):
            if trace is not None:
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                trace.append((label, e))

        mark("start")
        for l in range(L):
            W, b = params[2 * l], params[2 * l + 1]
            o = mod.sizes[l + 1]
            last = l == L - 1
            wb = mod._weight_bf16(l, W)
            wbs.append(wb)
            mark(f"l{l}.cast")
            if o % (64 * P) == 0 and use_nvls:
                # NVLS form: plain (2-CTA) GEMM into the own heap, then every rank pulls the switch-reduced
                # sum of its column block.  Row chunks pipeline the two: chunk c's reduction runs on the
                # side stream while the tensor cores are already on chunk c+1.
                w = o // P
                poff = plan.part_off(l & 1)
                part = sc.tensor(poff, (B, o), torch.bfloat16)
                nxt = torch.empty((B, w), device=x.device, dtype=torch.bfloat16)
                nch = max(1, min(mod.nvls_chunks, B // 512))
                bounds = [B * c // nch // 128 * 128 for c in range(nch)] + [B]
                side = mod._side_stream()
                main = torch.cuda.current_stream()
                if nch > 1:
                    nxt.record_stream(side)
                for c in range(nch):
                    r0, r1 = bounds[c], bounds[c + 1]
                    gemm_bf16(a[r0:r1], wb, bias=b, out=part[r0:r1])
                    if nch > 1:
                        side.wait_event(main.record_event())
                    with torch.cuda.stream(side if nch > 1 else main):
                        _native.check(lib.tadnn_reduce_cols_nvls(
                            sc.handle, C.c_ulonglong(poff + r0 * o * 2), r1 - r0, w, C.c_float(1.0), None, 0 if last else 1,
                            C.c_void_p(nxt[r0:].data_ptr()), C.c_ulonglong(w), int(last), 0, _symm.CH_TP,
                            C.c_void_p(_native.stream_ptr())), "tadnn_reduce_cols_nvls")
                if nch > 1:
                    main.wait_event(side.record_event())
                if last:
                    out = mod._finish(part)  # the owners multimem.st their column blocks into every rank's buffer
                a = nxt
            elif o % (64 * P) == 0:
                w = o // P
                push = [sc.heap_ptrs[q] + plan.slots_off + r * B * w * 2 for q in range(P)]  # slot r of rank q: [B, w]
                # every rank starts its tile sweep at a different owner's column block: at any moment the
                # P ranks push to P different destinations instead of all hammering rank 0's ingress
                nblk = (o + 255) // 256
                gemm_bf16(a, wb, bias=b, push=push, push_ldc=w, rs_width=w, n_rotate=((r + 1) % P) * nblk // P)
                mark(f"l{l}.gemm+push")
                nxt = torch.empty((B, w), device=x.device, dtype=torch.bfloat16)
                _native.check(lib.tadnn_reduce_slots(
                    sc.handle, C.c_ulonglong(plan.slots_off), B, w, C.c_float(1.0), None, 0 if last else 1,
                    C.c_void_p(nxt.data_ptr()), C.c_ulonglong(w), C.c_ulonglong(plan.gfull_off), int(last), 0, 0,
                    _symm.CH_TP, st), "tadnn_reduce_slots")
                if last:
                    out = mod._finish(sc.tensor(plan.gfull_off, (B, o), torch.bfloat16))
                a = nxt
            else:
                part = plan.small[:B * o].view(B, o)
                if _smalln_ok(o, a, "fwd"):
                    _smalln.linear_smalln_fwd(a, wb, bias=b, out=part)  # classifier-sized layer: GEMV-style kernel (K11)
                else:
                    gemm_bf16(a, wb, bias=b, out=part, out_dtype=torch.float32, tile_n=64)
                if P > 1:
                    sc.allreduce_(part, 1.0, channel=_symm.CH_TP)
                y = part.clone()
                if last:
                    out = y
                else:
                    assert o % (8 * P) == 0, f"hidden width {o} must be a multiple of 8*P"
                    a = torch.tanh(y[:, r * (o // P):(r + 1) * (o // P)]).to(torch.bfloat16).contiguous()
            mark(f"l{l}.reduce")
            if not last:
                acts.append(a)
        ctx.mod, ctx.B = mod, B
        ctx.save_for_backward(*acts, *wbs)
        return out

    @staticmethod
    def backward(ctx, gy):
        from ..ops.cnn import auto_split_k, colsum_bf16
        from ..ops.gemm import gemm_bf16
        mod, B = ctx.mod, ctx.B
        sc = mod._ctx()
        P, r = sc.world, sc.rank
        L = mod.num_layers
        plan = mod._plan(B)
        saved = ctx.saved_tensors
        acts, wbs = saved[:L], saved[L:]
        grads: List[Optional[torch.Tensor]] = [None] * (2 * L)
        dev = gy.device
        # replicated gradient of the last layer's output, bf16, pitch padded to a multiple of 8
        oL = mod.sizes[L]
        pitch = (oL + 7) // 8 * 8
        g_full = sc.tensor(plan.gfull_off, (B, pitch), torch.bfloat16)
        if pitch != oL:
            g_full.zero_()
        g_full[:, :oL].copy_(gy)
        gx = None
        offs = [plan.gfull_off, plan.gfull2_off]
        cur = 0
        side = mod._side_stream()
        main = torch.cuda.current_stream()
        import os
        fused_gather = P > 1 and os.environ.get("TADNN_MP_FUSED_GATHER", "1") not in ("0", "false")
        if fused_gather:
            g3, gfl = plan.gather_buffers()
        wait = None  # (flag pointer, target, columns per source) when g_full arrives through pushing GEMM epilogues
        for l in range(L - 1, -1, -1):
            o, i_sh = mod.sizes[l + 1], mod.sizes[l] // P
            pitch = (o + 7) // 8 * 8
            gv = torch.as_strided(g_full, (B, o), (pitch, 1))  # [B, o] with row pitch `pitch`
            a = acts[l]
            need_dx = l > 0 or ctx.needs_input_grad[1]
            gathered = None
            wkw = {} if wait is None else dict(wait_flags=wait[0], wait_target=wait[1], wait_cols_per_src=wait[2], flag_stride=P)
            nxt_wait, nxt_full = None, None
            if need_dx:
                # g'_r [B, i/P] = (g · W_r) (⊙ tanh' of the layer input if that input is a tanh output)
                tn = 256 if i_sh >= 256 else 0  # measured: 256-wide tiles beat 128-wide ones even at 1.7 waves
                if _smalln_ok(o, a, "dgrad"):
                    gr = _smalln.linear_smalln_dgrad(gv, wbs[l], aux_tanh_out=a if l > 0 else None)
                elif fused_gather and l > 0 and i_sh % 64 == 0:
                    # K13: this GEMM's epilogue pushes every finished 32x64 slab of its slice into the gathered buffer
                    # of ALL ranks and bumps their arrival counters; the GEMMs of layer l-1 consume the blocks as they
                    # land (flag waits in their TMA producer).  No all-gather kernel, no rank barrier.
                    d = (l - 1) % 3
                    i_full = i_sh * P
                    own = sc.tensor(g3[d], (B, i_full), torch.bfloat16)
                    plan.gtarget[d] += i_sh // 64
                    gemm_bf16(gv, wbs[l], b_mn_major=True, aux_tanh_out=a, tile_n=tn,
                              out=own[:, r * i_sh:(r + 1) * i_sh],
                              push=[sc.heap_ptrs[q] + g3[d] for q in range(P) if q != r], push_ldc=i_full, push_col_offset=r * i_sh,
                              sig_flags=[sc.heap_ptrs[q] + gfl[d] for q in range(P)], sig_src=r,
                              **(wkw or dict(flag_stride=P)))
                    gr = None
                    nxt_full = own
                    nxt_wait = (sc.heap_ptrs[r] + gfl[d], plan.gtarget[d], i_sh)
                else:
                    gr = gemm_bf16(gv, wbs[l], b_mn_major=True, aux_tanh_out=a if l > 0 else None, tile_n=tn, **wkw)
                if P > 1 and gr is not None:
                    # all-gather of the slice gradients on a side stream, into the OTHER buffer, so that it
                    # overlaps this layer's wgrad GEMM (the gather kernel needs few registers and co-resides)
                    side.wait_event(main.record_event())
                    gr.record_stream(side)
                    with torch.cuda.stream(side):
                        sc.allgather_cols_offset(gr, offs[cur ^ 1], _symm.CH_TP, side)
                        gathered = side.record_event()
            # dW_r [o, i/P] = gᵀ · a_r   (A: g stored [K=B, M=o];  B: a_r stored [K=B, N=i/P])
            if _smalln_ok(o, a, "wgrad"):
                dW, db = _smalln.linear_smalln_wgrad(gv, a)  # weight and bias gradient in one pass over a_r
            else:
                tile_n = 256 if i_sh > 128 else (128 if i_sh > 64 else 64)
                dW = gemm_bf16(gv, a, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32,
                               split_k=auto_split_k(o, i_sh, B, tile_n), tile_n=tile_n, **wkw)
                db = torch.zeros(o, device=dev, dtype=torch.float32)
                colsum_bf16(gv, db)  # (runs after the wgrad GEMM, whose watcher has seen every block of g arrive)
            grads[2 * l], grads[2 * l + 1] = dW, db
            if not need_dx:
                break
            if nxt_full is not None:
                g_full, wait = nxt_full, nxt_wait
            elif P > 1:
                main.wait_event(gathered)
                cur ^= 1
                g_full = sc.tensor(offs[cur], (B, i_sh * P), torch.bfloat16)
                wait = None
            else:
                g_full, wait = gr, None
            if l == 0:
                gx = g_full.to(gy.dtype).clone()
        return (None, gx, *grads)


class FusedMPMLP(torch.nn.Module):
    """``sizes = [i, h1, ..., o]``; tanh between layers; parameters in the reference layout."""

    def __init__(self, sizes: List[int], mpi=None, init: str = "reference"):
        super().__init__()
        h = mpi if mpi is not None else _mpi.current()
        self._mpi = mpi
        P = h.size()
        self.sizes = [int(s) for s in sizes]
        self.num_layers = len(sizes) - 1
        for s in self.sizes[:-1]:
            assert s % (8 * P) == 0, f"layer input width {s} must be a multiple of 8*P={8 * P}"
        self.weights = torch.nn.ParameterList()
        self.biases = torch.nn.ParameterList()
        for l in range(self.num_layers):
            i_sh, o = self.sizes[l] // P, self.sizes[l + 1]
            fan = i_sh if init == "reference" else self.sizes[l]
            s = 1.0 / math.sqrt(fan)
            self.weights.append(torch.nn.Parameter(torch.empty(o, i_sh).uniform_(-s, s)))
            self.biases.append(torch.nn.Parameter(torch.empty(o).uniform_(-s, s)))
        self._plans = {}

    @classmethod
    def from_sequential(cls, model, mpi=None):
        """Share the parameters of an ``MPInitialLinear / MPTanh / MPBaseLinear`` Sequential."""
        from ..nn import mp
        lins = [m for m in model.children() if isinstance(m, mp._MPLinear)]
        assert lins, "no MP linear layers found"
        kinds = [type(m).__name__ for m in model.children() if not isinstance(m, mp._MPReshape)]
        for k, name in enumerate(kinds):
            assert name == ("MPTanh" if k % 2 else ("MPInitialLinear" if k == 0 else "MPBaseLinear")), \
                f"unsupported MP stack {kinds}"
        sizes = [lins[0].in_features] + [m.out_features for m in lins]
        self = cls.__new__(cls)
        torch.nn.Module.__init__(self)
        self._mpi = mpi if mpi is not None else lins[0]._mpi
        self.sizes, self.num_layers = sizes, len(lins)
        self.weights = torch.nn.ParameterList([m.weight for m in lins])
        self.biases = torch.nn.ParameterList([m.bias for m in lins])
        self._plans = {}
        return self

    # ------------------------------------------------------------------
    # How the K-split partial products of a wide layer are summed across ranks:
    #   "push": the GEMM epilogue TMA-stores each column block into its owner's staging slot over NVLink,
    #           reduce_slots sums the slots locally (works on any P2P-capable box);
    #   "nvls": the GEMM writes locally, the owner pulls the sum through the switch (multimem.ld_reduce),
    #           optionally pipelined over `nvls_chunks` row chunks.  Needs the VMM + multicast heap.
    @property
    def reduce_mode(self) -> str:
        import os
        return getattr(self, "_reduce_mode", None) or os.environ.get("TADNN_MP_REDUCE", "push")

    @reduce_mode.setter
    def reduce_mode(self, v: str) -> None:
        assert v in ("push", "nvls")
        self._reduce_mode = v

    @property
    def nvls_chunks(self) -> int:
        import os
        return int(getattr(self, "_nvls_chunks", 0) or os.environ.get("TADNN_MP_NVLS_CHUNKS", "2"))

    @nvls_chunks.setter
    def nvls_chunks(self, v: int) -> None:
        self._nvls_chunks = int(v)

    # dtype of the returned activations: fp32 (default, what a PyTorch loss expects) or the bf16 the last
    # layer produced (half the bytes; pair it with ops.loss.mse_fused on very wide outputs)
    out_bf16 = False

    def _finish(self, t: torch.Tensor) -> torch.Tensor:
        """Copy the last layer's result out of the (re-used) symmetric buffer."""
        return t.clone() if self.out_bf16 else t.float()

    def _handle(self):
        return self._mpi if self._mpi is not None else _mpi.current()

    def _ctx(self) -> _symm.SymmContext:
        sc = getattr(self._handle(), "symm", None)
        if sc is None:
            raise _native.NativeError("FusedMPMLP needs the symmetric-memory context (mpi.start(True) on CUDA)")
        return sc

    def _side_stream(self):
        st = getattr(self, "_side", None)
        if st is None:
            st = self._side = torch.cuda.Stream()
        return st

    def _plan(self, B: int) -> _Plan:
        p = self._plans.get(B)
        if p is None:
            p = self._plans[B] = _Plan(self._ctx(), B, self.sizes)
        return p

    def _weight_bf16(self, l: int, W: torch.Tensor) -> torch.Tensor:
        """bf16 operand copy of layer l's weight: kept across steps and refreshed by ``sgd_step`` in the same pass
        as the update; re-cast only when the fp32 master was changed behind its back (tensor version bump)."""
        cache = self.__dict__.setdefault("_wb_cache", {})
        ent = cache.get(l)
        if ent is not None and ent[1] == W._version and ent[2] == W.data_ptr():
            return ent[0]
        wb = _cast_bf16(W)
        cache[l] = (wb, W._version, W.data_ptr())
        return wb

    @torch.no_grad()
    def sgd_step(self, lr: float, weight_decay: float = 0.0) -> None:
        """In-place SGD on every weight/bias from ``.grad`` (then dropped), fp32 master and bf16 operand copy in one
        kernel per weight (csrc/cnn.cu: tadnn_sgd_step_shadow).  The reference's in-place step: datamodule.lua:142."""
        lib = _native.lib()
        st = C.c_void_p(_native.stream_ptr())
        cache = self.__dict__.setdefault("_wb_cache", {})
        for l, (W, b) in enumerate(zip(self.weights, self.biases)):
            if W.grad is not None:
                g = W.grad if W.grad.is_contiguous() else W.grad.contiguous()
                ent = cache.get(l)
                fresh = ent is not None and ent[1] == W._version and ent[2] == W.data_ptr()
                if W.is_cuda and W.numel() % 4 == 0 and W.is_contiguous() and g.dtype == torch.float32:
                    wb = ent[0] if fresh else torch.empty(W.shape, device=W.device, dtype=torch.bfloat16)
                    _native.check(lib.tadnn_sgd_step_shadow(C.c_void_p(W.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(wb.data_ptr()),
                                                            C.c_longlong(W.numel()), C.c_float(lr), C.c_float(weight_decay), st),
                                  "tadnn_sgd_step_shadow")
                    cache[l] = (wb, W._version, W.data_ptr())
                else:
                    W.add_(g + weight_decay * W if weight_decay else g, alpha=-lr)
                W.grad = None
            if b.grad is not None:
                b.add_(b.grad + weight_decay * b if weight_decay else b.grad, alpha=-lr)
                b.grad = None

    def _params(self):
        out = []
        for w, b in zip(self.weights, self.biases):
            out += [w, b]
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x2 = x.reshape(x.shape[0], -1) if x.dim() > 2 else (x.reshape(1, -1) if x.dim() == 1 else x)
        assert x2.shape[1] == self.sizes[0], f"expected {self.sizes[0]} input features, got {x2.shape[1]}"
        h = self._handle()
        if x2.is_cuda and getattr(h, "symm", None) is not None and _native.available():
            return _Fn.apply(self, x2.contiguous(), *self._params())
        return self.reference_forward(x2)

    def reference_forward(self, x2: torch.Tensor) -> torch.Tensor:
        """Plain PyTorch statement (per-op MP semantics: narrow, linear, sum-all-reduce, tanh)."""
        from ..nn.mp import _AllReduceSum, _Narrow
        h = self._handle()
        a = x2.float()
        for l in range(self.num_layers):
            xs = _Narrow.apply(a, h, False) if (a.requires_grad and h.size() > 1) else \
                a.narrow(1, h.rank() * (a.shape[1] // h.size()), a.shape[1] // h.size())
            y = F.linear(xs, self.weights[l].float(), self.biases[l].float())
            if h.size() > 1:
                y = _AllReduceSum.apply(y, h)
            a = torch.tanh(y) if l < self.num_layers - 1 else y
        return a
