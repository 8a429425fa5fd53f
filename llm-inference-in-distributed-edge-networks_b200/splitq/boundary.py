"""Host side of the boundary path: thin, allocation-aware wrappers over the C ABI.

Stage order (BASELINE.json north_star): importance scoring -> importance-ranked bit-width
selection -> absmax/quantise/bit-pack into one payload -> transfer -> dequantise (optionally
fused with the downstream half's first norm).  Every function here takes and returns CUDA
tensors; the arithmetic happens in lib/libsplitq.so.  Reference lines each stage replaces are
cited in include/splitq.h.
"""
from __future__ import annotations

import ctypes as C
import threading
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from . import _native as N

Ladder = Sequence[Tuple[int, int]]  # [(token_count, bits), ...] least-important first


def _stream(t: torch.Tensor) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _ptr(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _need_cuda_f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if not (isinstance(t, torch.Tensor) and t.is_cuda):
        raise N.SplitqError(f"{name}: expected a CUDA tensor (there is no CPU path)")
    if t.dtype != torch.float32:
        raise N.SplitqError(f"{name}: expected float32, got {t.dtype}")
    return t.contiguous()


class _Workspace:
    """Grow-only scratch buffers so steady-state calls never allocate.  Kernel scratch is live from
    the launch until the kernel retires, so two calls may share a buffer only if they are ordered on
    one stream: the module-level instance keys its buffers by (device, tag, CURRENT STREAM), and every
    BoundaryPlan owns a private instance, so plans driven concurrently on different streams (or by
    different threads) never meet in the same scratch."""

    def __init__(self, per_stream: bool = True) -> None:
        self._bufs = {}
        self._per_stream = per_stream
        self._lock = threading.Lock()

    def get(self, device: torch.device, tag: str, nbytes: int) -> torch.Tensor:
        stream = torch.cuda.current_stream(device).cuda_stream if self._per_stream else 0
        key = (device.index, tag, stream)
        with self._lock:
            buf = self._bufs.get(key)
            if buf is None or buf.numel() < nbytes:
                buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
                self._bufs[key] = buf
        return buf


_WS = _Workspace()


def ladder_for_ratio(ratio: float, S: int, bits: int = 4) -> List[Tuple[int, int]]:
    """The reference's single rung: int(ratio*S) least-important tokens at 4 bits
    (pythia_model.py:169-170,179)."""
    k = int(ratio * S)
    return [(k, bits)] if k > 0 else []


def ladder_from_fractions(fracs: Sequence[Tuple[float, int]], S: int) -> List[Tuple[int, int]]:
    """[(0.25, 2), (0.25, 4), (0.25, 8)] -> lowest quarter 2-bit, next 4-bit, next 8-bit."""
    return [(int(f * S), b) for f, b in fracs]


def _ladder_arrays(ladder: Ladder):
    ks = [int(k) for k, _ in ladder]
    bs = [int(b) for _, b in ladder]
    return N.int_array(ks), N.int_array(bs), len(ks)


def payload_bytes(S: int, D: int, ladder: Ladder, scale_mode: int) -> int:
    ks, bs, n = _ladder_arrays(ladder)
    nbytes = N.load().splitq_payload_bytes(S, D, ks, bs, n, scale_mode)
    if nbytes == 0:
        raise N.SplitqError(f"payload_bytes: invalid arguments: {N.load().splitq_last_error().decode()}")
    return nbytes


def importance_from_attention(A: torch.Tensor, head_w: Optional[Sequence[float]] = None, want_col: bool = True,
                              want_last: bool = False, causal_hint: bool = False,
                              out_col: Optional[torch.Tensor] = None, out_last: Optional[torch.Tensor] = None,
                              workspace: Optional[_Workspace] = None):
    """A: (B,H,Sq,Sk) attention probabilities -> (col (B,Sk) or None, last (B,Sk) or None)."""
    A = _need_cuda_f32(A, "A")
    if A.dim() != 4:
        raise N.SplitqError(f"A: expected (B,H,Sq,Sk), got {tuple(A.shape)}")
    B, H, Sq, Sk = A.shape
    lib = N.load()
    if want_col and out_col is None:
        out_col = torch.empty(B, Sk, dtype=torch.float32, device=A.device)
    if want_last and out_last is None:
        out_last = torch.empty(B, Sk, dtype=torch.float32, device=A.device)
    ws_bytes = lib.splitq_importance_colmean_workspace_bytes(B, H, Sq, Sk) if want_col else 0
    ws = (workspace or _WS).get(A.device, "colmean", ws_bytes) if ws_bytes else None
    hw = None
    if head_w is not None:
        hw = N.float_array([float(x) for x in head_w])
        if len(hw) != H:
            raise N.SplitqError(f"head_w has {len(hw)} entries for {H} heads")
    with torch.cuda.device(A.device):
        N.check(lib.splitq_importance_colmean(_ptr(A), B, H, Sq, Sk, hw, _ptr(out_col if want_col else None),
                                              _ptr(out_last if want_last else None), _ptr(ws), ws_bytes,
                                              N.FLAG_CAUSAL_HINT if causal_hint else 0, _stream(A)))
    return (out_col if want_col else None), (out_last if want_last else None)


def importance_fused(q: torch.Tensor, k: torch.Tensor, scaling: float, causal: bool = True,
                     mode: int = N.FUSED_F16X3, head_w: Optional[Sequence[float]] = None, want_col: bool = True,
                     want_last: bool = False, out_col: Optional[torch.Tensor] = None,
                     out_last: Optional[torch.Tensor] = None, workspace: Optional[_Workspace] = None,
                     head_w2: Optional[Sequence[float]] = None):
    """q: (B,H,S,64), k: (B,KVH,S,64) post-RoPE fp32 CUDA tensors (any batch/head/seq strides, e.g.
    HF's `.view(...).transpose(1, 2)`) -> (col (B,S) or None, last (B,S) or None) without ever
    materialising the (S,S) attention matrix."""
    for name, t in (("q", q), ("k", k)):
        if not (t.is_cuda and t.dtype == torch.float32 and t.dim() == 4 and t.stride(3) == 1):
            raise N.SplitqError(f"{name}: expected a 4-d float32 CUDA tensor with contiguous head_dim")
    B, H, S, hd = q.shape
    KVH = k.shape[1]
    lib = N.load()
    if want_col and out_col is None:
        out_col = torch.empty(B, S, dtype=torch.float32, device=q.device)
    if want_last and out_last is None:
        out_last = torch.empty(B, S, dtype=torch.float32, device=q.device)
    ws_bytes = lib.splitq_importance_fused_workspace_bytes(B, H, KVH, S, hd)
    if ws_bytes == 0:
        raise N.SplitqError(f"importance_fused: unsupported shape q={tuple(q.shape)} k={tuple(k.shape)} (head_dim 64 only)")
    ws = (workspace or _WS).get(q.device, "fused", ws_bytes)
    hw = None
    if head_w is not None:
        hw = N.float_array([float(x) for x in head_w])
        if len(hw) != H:  # the C side reads exactly H host floats
            raise N.SplitqError(f"head_w has {len(hw)} entries for {H} heads")
    qs = (C.c_longlong * 3)(q.stride(0), q.stride(1), q.stride(2))
    ks = (C.c_longlong * 3)(k.stride(0), k.stride(1), k.stride(2))
    if head_w2 is not None:
        # a second weighting of the same per-head sums in the same pass -> third return value (B,S)
        hw2 = N.float_array([float(x) for x in head_w2])
        if len(hw2) != H:
            raise N.SplitqError(f"head_w2 has {len(hw2)} entries for {H} heads")
        out_col2 = torch.empty(B, S, dtype=torch.float32, device=q.device)
        with torch.cuda.device(q.device):
            N.check(lib.splitq_importance_fused_dual(_ptr(q), qs, _ptr(k), ks, B, H, KVH, S, hd, float(scaling), int(causal),
                                                     mode, hw, _ptr(out_col if want_col else None),
                                                     _ptr(out_last if want_last else None), hw2, _ptr(out_col2), _ptr(ws),
                                                     ws_bytes, _stream(q)))
        return (out_col if want_col else None), (out_last if want_last else None), out_col2
    with torch.cuda.device(q.device):
        N.check(lib.splitq_importance_fused(_ptr(q), qs, _ptr(k), ks, B, H, KVH, S, hd, float(scaling), int(causal),
                                            mode, hw, _ptr(out_col if want_col else None),
                                            _ptr(out_last if want_last else None), _ptr(ws), ws_bytes, _stream(q)))
    return (out_col if want_col else None), (out_last if want_last else None)


def select_bits(imp: torch.Tensor, ladder: Ladder, D: int, bits_out: Optional[torch.Tensor] = None,
                row_off_out: Optional[torch.Tensor] = None):
    """imp: (B,S) or (S,) -> (bits u8 (B,S), row_off u32-as-int32 (B,S))."""
    imp = _need_cuda_f32(imp, "importance")
    if imp.dim() == 1:
        imp = imp.unsqueeze(0)
    B, S = imp.shape
    if bits_out is None:
        bits_out = torch.empty(B, S, dtype=torch.uint8, device=imp.device)
    if row_off_out is None:
        row_off_out = torch.empty(B, S, dtype=torch.int32, device=imp.device)
    ks, bs, n = _ladder_arrays(ladder)
    with torch.cuda.device(imp.device):
        N.check(N.load().splitq_select_bits(_ptr(imp), B, S, D, ks, bs, n, _ptr(bits_out), _ptr(row_off_out),
                                            _stream(imp)))
    return bits_out, row_off_out


def select_bits_per_window(imp: torch.Tensor, bits: int, D: int, counts: Optional[Sequence[int]] = None,
                           mass_thresholds: Optional[Sequence[float]] = None, want_counts: bool = False):
    """One rung of `bits` bits whose size differs per window of imp (B,S): either `counts[b]` least
    important tokens (batched_quantization_helper's rule), or "top-rho": everything outside the
    fewest most-important tokens whose importance mass reaches mass_thresholds[b]
    (batched_top_rho_quantization, pythia_model.py:93-109).  -> (bits u8 (B,S), row_off (B,S)
    [, quantised-token counts int32 (B,) on the device])."""
    imp = _need_cuda_f32(imp, "importance")
    if imp.dim() == 1:
        imp = imp.unsqueeze(0)
    B, S = imp.shape
    if (counts is None) == (mass_thresholds is None):
        raise N.SplitqError("select_bits_per_window: give either counts or mass_thresholds")
    n = len(counts if counts is not None else mass_thresholds)
    if n != B:
        raise N.SplitqError(f"select_bits_per_window: {n} per-window values for {B} windows")
    bits_out = torch.empty(B, S, dtype=torch.uint8, device=imp.device)
    row_off_out = torch.empty(B, S, dtype=torch.int32, device=imp.device)
    k_out = torch.empty(B, dtype=torch.int32, device=imp.device) if want_counts else None
    ks = N.int_array([int(c) for c in counts]) if counts is not None else None
    ms = (C.c_double * B)(*[float(t) for t in mass_thresholds]) if mass_thresholds is not None else None
    with torch.cuda.device(imp.device):
        N.check(N.load().splitq_select_bits_per_window(_ptr(imp), B, S, D, int(bits), ks, ms, _ptr(bits_out),
                                                       _ptr(row_off_out), _ptr(k_out), _stream(imp)))
    return (bits_out, row_off_out, k_out) if want_counts else (bits_out, row_off_out)


def _check_pack_args(who: str, h: torch.Tensor, bits: torch.Tensor, row_off: torch.Tensor, payload: torch.Tensor,
                     nbytes: Optional[int]) -> None:
    """The C side trusts these pointers: a wrong dtype, device, layout or size is an out-of-bounds
    access there, so it is refused here."""
    B, S, _ = h.shape
    for name, t, dt in (("bits", bits, torch.uint8), ("row_off", row_off, torch.int32)):
        if not (isinstance(t, torch.Tensor) and t.device == h.device and t.dtype == dt and t.is_contiguous()
                and tuple(t.shape) == (B, S)):
            raise N.SplitqError(f"{who}: {name} must be a contiguous ({B},{S}) {dt} tensor on {h.device}")
    if not (payload.device == h.device and payload.dtype == torch.uint8 and payload.dim() == 2
            and payload.shape[0] == B and payload.stride(1) == 1 and payload.stride(0) % 16 == 0):
        raise N.SplitqError(f"{who}: payload must be a ({B}, nbytes) uint8 tensor on {h.device} with a row stride "
                            "that is a multiple of 16")
    if nbytes is not None and payload.shape[1] < nbytes:
        raise N.SplitqError(f"{who}: payload rows hold {payload.shape[1]} bytes, {nbytes} needed")


def quant_pack(h: torch.Tensor, bits: torch.Tensor, row_off: torch.Tensor, scale_mode: int, nbytes: int,
               payload: Optional[torch.Tensor] = None, cuda_aten_dequant: bool = False,
               workspace: Optional[_Workspace] = None) -> torch.Tensor:
    """h: (B,S,D) -> payload uint8 (B, nbytes).  `nbytes` = payload_bytes(S, D, ladder, scale_mode).
    cuda_aten_dequant marks the payload so that the receiver dequantises with q * fl(1/L) -- what the
    reference's `quantized / 7 * max_val` evaluates to when it runs on a GPU -- instead of the
    correctly rounded quotient CPU ATen computes (the default)."""
    h = _need_cuda_f32(h, "hidden_states")
    if h.dim() == 2:
        h = h.unsqueeze(0)
    B, S, D = h.shape
    if payload is None:
        payload = torch.empty(B, nbytes, dtype=torch.uint8, device=h.device)
    _check_pack_args("quant_pack", h, bits, row_off, payload, nbytes)
    lib = N.load()
    ws_bytes = lib.splitq_quant_pack_workspace_bytes(B, S, D, scale_mode)
    ws = (workspace or _WS).get(h.device, "quant", ws_bytes) if ws_bytes else None
    with torch.cuda.device(h.device):
        N.check(lib.splitq_quant_pack(_ptr(h), B, S, D, _ptr(bits), _ptr(row_off),
                                      scale_mode | (N.FLAG_DEQUANT_RECIP if cuda_aten_dequant else 0), _ptr(payload),
                                      payload.stride(0), _ptr(ws), ws_bytes, _stream(h)))
    return payload


def quant_pack_group_noalloc(h: torch.Tensor, bits: torch.Tensor, row_off: torch.Tensor,
                             payload: torch.Tensor) -> torch.Tensor:
    """Group-mode quantise + pack for callers that may not allocate on the device (a runtime that
    reserved all of HBM for itself, or a hook that must not allocate while enqueuing).  The
    class-maximum scratch is a fixed array inside the library, so only inputs and the output
    `payload` (B, nbytes) are passed -- all of them existing tensors; nothing is allocated here.
    At most _native.NOALLOC_MAX_WINDOWS windows per call.  Calls share that array: do not let two
    of them overlap on the GPU.  Make the first call on a device before any stream capture."""
    if not (h.is_cuda and h.dtype == torch.float32 and h.is_contiguous() and h.dim() == 3):
        raise N.SplitqError("quant_pack_group_noalloc: h must be a contiguous (B,S,D) float32 CUDA tensor")
    B, S, D = h.shape
    _check_pack_args("quant_pack_group_noalloc", h, bits, row_off, payload, None)
    with torch.cuda.device(h.device):
        N.check(N.load().splitq_quant_pack_group_noalloc(_ptr(h), B, S, D, _ptr(bits), _ptr(row_off), _ptr(payload),
                                                         payload.stride(0), _stream(h)))
    return payload


def dequant(payload: torch.Tensor, S: int, D: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """payload (B, nbytes) uint8 -> (B,S,D) fp32."""
    B = payload.shape[0]
    if out is None:
        out = torch.empty(B, S, D, dtype=torch.float32, device=payload.device)
    with torch.cuda.device(payload.device):
        N.check(N.load().splitq_dequant(_ptr(payload), payload.stride(0), B, S, D, _ptr(out), _stream(payload)))
    return out


def dequant_norm(payload: torch.Tensor, S: int, D: int, kind: int, gamma: torch.Tensor,
                 beta: Optional[torch.Tensor], eps: float, gamma2: Optional[torch.Tensor] = None,
                 beta2: Optional[torch.Tensor] = None, want_resid: bool = True, outs=None):
    """-> (resid or None, normed, normed2 or None), each (B,S,D) fp32."""
    B = payload.shape[0]
    dev = payload.device
    if outs is None:
        mk = lambda: torch.empty(B, S, D, dtype=torch.float32, device=dev)  # noqa: E731
        outs = (mk() if want_resid else None, mk(), mk() if gamma2 is not None else None)
    resid, normed, normed2 = outs
    with torch.cuda.device(dev):
        N.check(N.load().splitq_dequant_norm(_ptr(payload), payload.stride(0), B, S, D, kind, _ptr(gamma),
                                             _ptr(beta), float(eps), _ptr(gamma2), _ptr(beta2), _ptr(resid),
                                             _ptr(normed), _ptr(normed2), _stream(payload)))
    return resid, normed, normed2


def channel_fake_quant(h: torch.Tensor, method: str, out: Optional[torch.Tensor] = None, want_codes: bool = False):
    """Channel-wise quantise->dequantise of (B,S,D) hidden states; method in {"channel_8",
    "channel_4", "channel_1_mean", "channel_1_max"} (qwen_layer_wise.py:106-152).
    Returns (h_out, per-channel scales (D,), int8 codes or None); `out=h` works in place."""
    if method not in N.CHANNEL_METHODS:
        raise ValueError(f"Unknown method: {method}")
    h = _need_cuda_f32(h, "hidden_states")
    if h.dim() == 2:
        h = h.unsqueeze(0)
    B, S, D = h.shape
    out = torch.empty_like(h) if out is None else out
    stats = torch.empty(D, dtype=torch.float32, device=h.device)
    codes = torch.empty(B, S, D, dtype=torch.int8, device=h.device) if want_codes else None
    lib = N.load()
    ws_bytes = lib.splitq_channel_workspace_bytes(B, S, D)
    ws = _WS.get(h.device, "channel", ws_bytes)
    with torch.cuda.device(h.device):
        N.check(lib.splitq_channel_fake_quant(_ptr(h), B, S, D, N.CHANNEL_METHODS[method], _ptr(stats), _ptr(out),
                                              _ptr(codes), _ptr(ws), ws_bytes, _stream(h)))
    return out, stats, codes


def channel_payload_bytes(S: int, D: int, method: str) -> int:
    if method not in N.CHANNEL_METHODS:
        raise ValueError(f"Unknown method: {method}")
    return N.load().splitq_channel_payload_bytes(S, D, N.CHANNEL_METHODS[method])


def channel_quant_pack(h: torch.Tensor, method: str, payload: Optional[torch.Tensor] = None,
                       workspace: Optional[_Workspace] = None) -> torch.Tensor:
    """The channel-wise quantisers as a real payload (SPQ1, scale mode SCALE_CHANNEL_BASE + method): per
    window the per-channel statistic and the 8 / 4 / 2-bit codes of every row.  `dequant` / `dequant_norm`
    read it back; for one window pack -> dequant equals channel_fake_quant bit for bit."""
    if method not in N.CHANNEL_METHODS:
        raise ValueError(f"Unknown method: {method}")
    h = _need_cuda_f32(h, "hidden_states")
    if h.dim() == 2:
        h = h.unsqueeze(0)
    B, S, D = h.shape
    nbytes = channel_payload_bytes(S, D, method)
    if nbytes == 0:
        raise N.SplitqError(f"channel_quant_pack: unsupported shape {tuple(h.shape)}")
    if payload is None:
        payload = torch.empty(B, nbytes, dtype=torch.uint8, device=h.device)
    if not (payload.device == h.device and payload.dtype == torch.uint8 and payload.shape[0] == B
            and payload.shape[1] >= nbytes and payload.stride(1) == 1 and payload.stride(0) % 16 == 0):
        raise N.SplitqError(f"channel_quant_pack: payload must be a ({B}, >= {nbytes}) uint8 tensor on {h.device}")
    lib = N.load()
    ws_bytes = lib.splitq_channel_workspace_bytes(B, S, D)
    ws = (workspace or _WS).get(h.device, "channel", ws_bytes)
    with torch.cuda.device(h.device):
        N.check(lib.splitq_channel_quant_pack(_ptr(h), B, S, D, N.CHANNEL_METHODS[method], _ptr(payload), payload.stride(0),
                                              _ptr(ws), ws_bytes, _stream(h)))
    return payload


def transfer(dst: torch.Tensor, src: torch.Tensor, nbytes: Optional[int] = None) -> torch.Tensor:
    """Payload hop between the halves' devices (cudaMemcpyPeerAsync; same device = D2D copy).
    Enqueued on the current stream of the SOURCE device."""
    nbytes = src.numel() * src.element_size() if nbytes is None else nbytes
    with torch.cuda.device(src.device):
        N.check(N.load().splitq_transfer_peer(_ptr(dst), dst.device.index, _ptr(src), src.device.index, nbytes,
                                              _stream(src)))
    return dst


@dataclass
class BoundaryResult:
    importance: torch.Tensor
    bits: torch.Tensor
    payload: torch.Tensor
    hidden: torch.Tensor


class BoundaryPlan:
    """Pre-sized buffers for repeated passes of the whole path over (B,H,S,S) attention maps and
    (B,S,D) hidden states; `run` allocates nothing."""

    def __init__(self, B: int, H: int, S: int, D: int, ladder: Ladder, scale_mode: int = N.SCALE_GROUP,
                 method: str = "regular_importance", src_device="cuda:0", dst_device=None,
                 causal_hint: bool = False, KVH: Optional[int] = None, fused_mode: int = N.FUSED_F16X3):
        self.B, self.H, self.S, self.D = B, H, S, D
        self.ladder, self.scale_mode, self.method, self.causal_hint = list(ladder), scale_mode, method, causal_hint
        self.src = torch.device(src_device)
        self.dst = torch.device(dst_device) if dst_device is not None else self.src
        self.nbytes = payload_bytes(S, D, self.ladder, scale_mode)
        self.transport = None
        self.KVH = KVH or H
        self.fused_mode = fused_mode
        f32 = dict(dtype=torch.float32, device=self.src)
        self.imp = torch.empty(B, S, **f32)
        self.bits = torch.empty(B, S, dtype=torch.uint8, device=self.src)
        self.row_off = torch.empty(B, S, dtype=torch.int32, device=self.src)
        self.payload_src = torch.empty(B, self.nbytes, dtype=torch.uint8, device=self.src)
        self.payload_dst = torch.empty(B, self.nbytes, dtype=torch.uint8, device=self.dst)
        self.h_out = torch.empty(B, S, D, dtype=torch.float32, device=self.dst)
        # private scratch (not shared with any other plan or stream); the small one is sized now, the
        # scoring scratch (materialised or fused, whichever this plan is driven with) on first use
        self.ws = _Workspace(per_stream=False)
        self.ws.get(self.src, "quant", N.load().splitq_quant_pack_workspace_bytes(B, S, D, scale_mode))
        self._dst_done = None  # two-device plans: the downstream device has finished reading payload_dst

    def score(self, A: torch.Tensor) -> torch.Tensor:
        last = self.method == "last_row"
        importance_from_attention(A, want_col=not last, want_last=last, causal_hint=self.causal_hint,
                                  out_col=self.imp, out_last=self.imp, workspace=self.ws)
        return self.imp

    def score_fused(self, q: torch.Tensor, k: torch.Tensor, scaling: Optional[float] = None) -> torch.Tensor:
        """Importance straight from the split layer's post-RoPE Q (B,H,S,64) and K (B,KVH,S,64)."""
        last = self.method == "last_row"
        scaling = q.shape[-1] ** -0.5 if scaling is None else scaling
        importance_fused(q, k, scaling, causal=True, mode=self.fused_mode, want_col=not last, want_last=last,
                         out_col=self.imp, out_last=self.imp, workspace=self.ws)
        return self.imp

    def send(self, h: torch.Tensor) -> torch.Tensor:
        select_bits(self.imp, self.ladder, self.D, self.bits, self.row_off)
        quant_pack(h, self.bits, self.row_off, self.scale_mode, self.nbytes, self.payload_src, workspace=self.ws)
        return self.payload_src

    def hop(self) -> torch.Tensor:
        if self.transport is not None:  # e.g. transport.PairExchange between ranks
            return self.transport.send(self.payload_src, self.payload_dst)
        if self.dst != self.src and self._dst_done is not None:
            # the copy is enqueued on the SOURCE device's stream but overwrites a buffer that the
            # DESTINATION device's stream may still be dequantising from (the previous pass)
            torch.cuda.current_stream(self.src).wait_event(self._dst_done)
        transfer(self.payload_dst, self.payload_src)
        if self.dst != self.src:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.src))
            torch.cuda.current_stream(self.dst).wait_event(ev)
        return self.payload_dst

    def receive(self) -> torch.Tensor:
        out = dequant(self.payload_dst, self.S, self.D, self.h_out)
        if self.dst != self.src:
            self._dst_done = torch.cuda.Event()
            self._dst_done.record(torch.cuda.current_stream(self.dst))
        return out

    def run(self, A: torch.Tensor, h: torch.Tensor) -> BoundaryResult:
        self.score(A)
        self.send(h)
        self.hop()
        self.receive()
        return BoundaryResult(self.imp, self.bits, self.payload_dst, self.h_out)

    def run_fused(self, q: torch.Tensor, k: torch.Tensor, h: torch.Tensor) -> BoundaryResult:
        self.score_fused(q, k)
        self.send(h)
        self.hop()
        self.receive()
        return BoundaryResult(self.imp, self.bits, self.payload_dst, self.h_out)

    def capture(self, *inputs: torch.Tensor, fused: bool = True) -> "torch.cuda.CUDAGraph":
        """Record one whole pass (score, select, quantise/pack, hop, dequantise) over these exact
        input tensors into a CUDA graph; `graph.replay()` then re-runs it with a single launch,
        which is what matters at the reference's own batch size of one window, where the handful of
        short kernels is launch-bound.  The entry points neither allocate nor synchronise, so they
        can be captured; one eager pass is run first because a few of them do one-time set-up on
        their first call (kernel attributes, driver entry points, workspace growth)."""
        run = (lambda: self.run_fused(*inputs)) if fused else (lambda: self.run(*inputs))
        run()
        torch.cuda.synchronize(self.src)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            run()
        return graph

    def run_host_fused(self, q_host, k_host, h_host, out_host) -> torch.Tensor:
        """End-to-end call with HOST buffers for the fused path: H2D of Q, K and the hidden states,
        the whole path, D2H of the dequantised hidden states.  Blocks until `out_host` is filled."""
        if getattr(self, "_q_dev", None) is None:
            self._q_dev = torch.empty(q_host.shape, dtype=torch.float32, device=self.src)
            self._k_dev = torch.empty(k_host.shape, dtype=torch.float32, device=self.src)
            self._h_dev = torch.empty(self.B, self.S, self.D, dtype=torch.float32, device=self.src)
        self._q_dev.copy_(q_host, non_blocking=True)
        self._k_dev.copy_(k_host, non_blocking=True)
        self._h_dev.copy_(h_host, non_blocking=True)
        self.run_fused(self._q_dev, self._k_dev, self._h_dev)
        with torch.cuda.device(self.dst):
            out_host.copy_(self.h_out, non_blocking=True)
            torch.cuda.current_stream(self.dst).synchronize()
        return out_host

    def run_host(self, A_host: torch.Tensor, h_host: torch.Tensor, out_host: torch.Tensor) -> torch.Tensor:
        """End-to-end call with HOST buffers (pinned for full speed): H2D of this step's attention
        maps and hidden states, the whole path, D2H of the dequantised hidden states.  Blocks until
        the result is in `out_host`."""
        if getattr(self, "_A_dev", None) is None:
            self._A_dev = torch.empty(self.B, self.H, self.S, self.S, dtype=torch.float32, device=self.src)
            self._h_dev = torch.empty(self.B, self.S, self.D, dtype=torch.float32, device=self.src)
        self._A_dev.copy_(A_host, non_blocking=True)
        self._h_dev.copy_(h_host, non_blocking=True)
        self.run(self._A_dev, self._h_dev)
        with torch.cuda.device(self.dst):
            out_host.copy_(self.h_out, non_blocking=True)
            torch.cuda.current_stream(self.dst).synchronize()
        return out_host

    def host_bytes_per_run(self, fused: bool = False):
        """(h2d, d2h) bytes moved by one run_host() / run_host_fused()."""
        if fused:
            return 4 * self.B * self.S * (64 * (self.H + self.KVH) + self.D), 4 * self.B * self.S * self.D
        return 4 * self.B * (self.H * self.S * self.S + self.S * self.D), 4 * self.B * self.S * self.D

    #: kernels launched by one run(): colsum partial + finalize, select, [memset+class absmax,] pack, dequant
    def launches_per_run(self, fused: bool = False) -> int:
        n = 5 if fused else (2 if self.method != "last_row" else 1)  # fused: split, pass 1, lse merge, pass 2, finalize
        n += 1
        n += 2 if self.scale_mode == N.SCALE_GROUP else 1
        n += 1
        return n


class HostPipeline:
    """Host-buffer front end that keeps PCIe and the GPU busy at the same time.

    `submit()` takes one step's inputs in (pinned) HOST memory, enqueues their host->device copy
    on a copy stream, the whole boundary path on the compute stream and the device->host read of
    the dequantised hidden states on a second copy stream, and returns without waiting; successive
    steps alternate between `depth` buffer sets, so the upload of step i+1 and the download of step
    i-1 overlap the kernels of step i.  `drain()` blocks until every submitted result is on the
    host.  Works for the fused (q, k, h) and the materialised (A, h) inputs."""

    def __init__(self, make_plan, depth: int = 2, fused: bool = True):
        self.plans = [make_plan() for _ in range(depth)]
        self.fused = fused
        dev = self.plans[0].src
        self.compute = torch.cuda.current_stream(dev)
        self.h2d = torch.cuda.Stream(dev)
        self.d2h = torch.cuda.Stream(self.plans[0].dst)
        self.slot = 0
        for p in self.plans:
            p._in = None
            p._compute_done = None
            p._d2h_done = None

    def submit(self, inputs_host: Sequence[torch.Tensor], out_host: torch.Tensor) -> None:
        p = self.plans[self.slot]
        self.slot = (self.slot + 1) % len(self.plans)
        if p._in is None:
            p._in = [torch.empty(t.shape, dtype=t.dtype, device=p.src) for t in inputs_host]
        with torch.cuda.stream(self.h2d):
            if p._compute_done is not None:  # the step that last used this buffer set has consumed its inputs
                self.h2d.wait_event(p._compute_done)
            for d, h in zip(p._in, inputs_host):
                d.copy_(h, non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(self.h2d)
        self.compute.wait_event(ready)
        if p._d2h_done is not None:  # ... and its result has left the device
            self.compute.wait_event(p._d2h_done)
        if self.fused:
            p.run_fused(*p._in)
        else:
            p.run(*p._in)
        p._compute_done = torch.cuda.Event()
        p._compute_done.record(torch.cuda.current_stream(p.dst))
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(p._compute_done)
            out_host.copy_(p.h_out, non_blocking=True)
            p._d2h_done = torch.cuda.Event()
            p._d2h_done.record(self.d2h)

    def drain(self) -> None:
        self.d2h.synchronize()
        self.compute.synchronize()


class OverlappedBoundary:
    """Device-resident variant of the same software pipeline: the scoring of step i+1 runs on the
    main stream while select + quantise/pack, the payload hop (peer copy, IPC pull or NCCL pair
    exchange) and the dequantise of step i run on a side stream; `depth` buffer sets alternate.
    With `pack_on_side=False` select + quantise/pack stay on the main stream behind the scoring
    (they then delay the next step's scoring by their ~50 us instead of running beside its
    memory-bound operand split).  The input tensors of a step are read on both streams: keep them
    alive and unmodified until `join()` / `drain()` or until the same slot comes round again."""

    def __init__(self, make_plan, depth: int = 2, fused: bool = True, pack_on_side: bool = True,
                 side_priority: int = 0, main_priority=None):
        """main_priority: None = score on the caller's current stream; an int = score on a stream of
        that CUDA priority (-1 and below: higher than default), which lets the block scheduler place
        the scoring kernels' CTAs ahead of the side stream's whenever an SM frees up.  The caller's
        current stream is joined at every step and can be re-joined with `join()`."""
        self.plans = [make_plan() for _ in range(depth)]
        self.fused = fused
        self.pack_on_side = pack_on_side
        dev = self.plans[0].src
        self.caller = torch.cuda.current_stream(dev)
        self.main = self.caller if main_priority is None else torch.cuda.Stream(dev, priority=main_priority)
        self.side = torch.cuda.Stream(dev, priority=side_priority)
        self.slot = 0
        for p in self.plans:
            p._xfer_done = None

    def _pack(self, p, h) -> None:
        tr = p.transport
        if hasattr(tr, "before_pack"):  # transport.IpcPairExchange: the partner has fetched this slot's last payload
            tr.before_pack(p)
        p.send(h)
        if hasattr(tr, "after_pack"):   # ... and may fetch the new one
            tr.after_pack(p)

    def step(self, *inputs: torch.Tensor, score_events=None, side_event=None) -> "BoundaryPlan":
        """score_events = (start, end) timing events recorded around the scoring on the main stream;
        side_event = an event recorded on the side stream when this step's dequantise is enqueued."""
        p = self.plans[self.slot]
        self.slot = (self.slot + 1) % len(self.plans)
        if self.main is not self.caller:  # inputs were produced on the caller's stream
            self.main.wait_stream(self.caller)
        if p._xfer_done is not None:  # this slot's importance, payload and output buffers are free again
            self.main.wait_event(p._xfer_done)
        if score_events is not None:
            score_events[0].record(self.main)
        with torch.cuda.stream(self.main):
            if self.fused:
                p.score_fused(inputs[0], inputs[1])
            else:
                p.score(inputs[0])
            if score_events is not None:
                score_events[1].record(self.main)
            if not self.pack_on_side:
                self._pack(p, inputs[-1])
        handoff = torch.cuda.Event()
        handoff.record(self.main)
        with torch.cuda.stream(self.side):
            self.side.wait_event(handoff)
            if self.pack_on_side:
                self._pack(p, inputs[-1])
            p.hop()
            p.receive()
            if p._dst_done is not None:  # two-device plan: the dequantise ran on the other device's stream
                self.side.wait_event(p._dst_done)
            p._xfer_done = torch.cuda.Event()
            p._xfer_done.record(self.side)
            if side_event is not None:
                side_event.record(self.side)
        return p

    def join(self) -> None:
        """Make the caller's stream wait for everything submitted so far (no host sync)."""
        self.main.wait_stream(self.side)
        if self.main is not self.caller:
            self.caller.wait_stream(self.main)

    def drain(self) -> None:
        self.side.synchronize()
        self.main.synchronize()
