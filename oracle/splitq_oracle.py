"""CPU oracle for the split-inference boundary path.  TEST INFRASTRUCTURE ONLY.

This file restates, on the CPU, the arithmetic of the reference's boundary path
(sv-goat/LLM-Inference-in-Distributed-Edge-Networks).  It exists so that the CUDA
kernels can be checked against it; it is never imported by the product package
(`splitq`), only by `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py`.

Pinning status
--------------
The reference ships no tests and no golden vectors.  The restatement is pinned by
running the reference's own Python functions (imported from /root/reference, never
copied) on seeded inputs in the build container: `oracle/make_golden.py` writes their
outputs to `tests/golden/*.npz`, and `tests/test_oracle_golden.py` replays this file
against those fixtures (importance: exact same ATen ops -> bit-equal; fake-quant block:
bit-equal; wrapper NLL: bit-equal).  Stages the reference does not have at all (integer
codes as a stored type, the packed payload, per-token scales, the 8/2-bit ladder,
LayerNorm/RMSNorm fused dequant) are *our* specification; for those this file IS the
spec and parity is "unpinned by the reference" by construction.

Arithmetic is done with torch CPU fp32 ops in the same order as the reference so that
codes come out bit-identical; byte packing is numpy.

Reference file:line citations are relative to /root/reference/Experiments/.
"""
from __future__ import annotations

import math
import struct
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

# --------------------------------------------------------------------------------------
# payload format constants (mirrored in include/splitq.h; the format is ours, see above)
# --------------------------------------------------------------------------------------
MAGIC = 0x31515053  # "SPQ1" little-endian
HEADER_BYTES = 32
SCALE_GROUP = 0  # one scale per bit-width class (reference: one scale for the selected set)
SCALE_TOKEN = 1  # one scale per token (north-star extension)
BITS_RAW = 32  # token crosses the boundary as raw fp32
GROUP_SLOTS = 4  # scale slot: 8-bit -> 0, 4-bit -> 1, 2-bit -> 2, slot 3 unused


def pad16(n: int) -> int:
    return (n + 15) & ~15


def levels(bits: int) -> int:
    """Symmetric positive level count L_b = 2^(b-1) - 1.

    4 bit -> 7 (pythia_model.py:179-182, qwen_layer_wise.py:63-64: num_levels/2 - 1),
    8 bit -> 127 and ternary -> 1 (qwen_layer_wise.py:108-111, :147-148)."""
    return (1 << (bits - 1)) - 1


def slot_of_bits(bits: int) -> int:
    return {8: 0, 4: 1, 2: 2}[bits]


def row_bytes(D: int, bits: int) -> int:
    return pad16((D * bits + 7) // 8)


# --------------------------------------------------------------------------------------
# (a1-a4) importance scoring over materialised attention
# --------------------------------------------------------------------------------------
def importance_regular(A: torch.Tensor) -> torch.Tensor:
    """Column-wise mean, heads first then query rows.  A: (1,H,S,S) fp32 -> (S,).
    Pythia-70M/last_row_exp.py:13-23, Qwen2-0.5B/main.py:46-56."""
    per_head_mean = torch.mean(A, dim=1)
    return torch.mean(per_head_mean, dim=1).squeeze(0)


def importance_last_row(A: torch.Tensor) -> torch.Tensor:
    """Mean over heads of the last query row.  last_row_exp.py:24-30, Qwen main.py:80-86."""
    return torch.mean(A[:, :, -1, :], dim=1).squeeze(0)


def importance_weighted(A: torch.Tensor, head_w: Sequence[float]) -> torch.Tensor:
    """Head-weighted column mean.  Qwen2-0.5B/main.py:57-78 (weights are per-head scalars)."""
    acc = torch.zeros_like(A[:, 0, :, :])
    for h in range(A.shape[1]):
        acc += A[:, h, :, :] * head_w[h]
    return torch.mean(acc, dim=1).squeeze(0)


def get_importance_order(method: str, attention_map, num_layers: int, head_weights=None) -> List[torch.Tensor]:
    """Same contract as last_row_exp.py:9-45 / Qwen main.py:21-98."""
    out: List[torch.Tensor] = []
    running = 0
    for l in range(num_layers):
        if method == "regular_importance":
            out.append(importance_regular(attention_map[l]))
        elif method == "weighted_importance":
            out.append(importance_weighted(attention_map[l], head_weights[l]))
        elif method == "last_row":
            out.append(importance_last_row(attention_map[l]))
        elif method == "aggregate_till":
            # last_row_exp.py:31-39: heads mean, squeeze, mean over rows, running mean over layers
            cur = torch.mean(torch.mean(attention_map[l], dim=1).squeeze(0), dim=0)
            running = running + cur
            out.append(running / (l + 1))
        else:
            raise ValueError(f"Unknown method: {method}")
    return out


def attention_probs(q: torch.Tensor, k: torch.Tensor, scaling: float, causal: bool = True) -> torch.Tensor:
    """HF eager attention probabilities (third-party: transformers
    models/qwen2/modeling_qwen2.py eager_attention_forward; reference call sites
    last_row_exp.py:94, Qwen main.py:159).  q: (B,H,S,hd), k: (B,KVH,S,hd) -> (B,H,S,S) fp32."""
    B, H, S, hd = q.shape
    rep = H // k.shape[1]
    kk = k[:, :, None, :, :].expand(B, k.shape[1], rep, S, hd).reshape(B, H, S, hd)
    s = torch.matmul(q, kk.transpose(2, 3)) * scaling
    if causal:
        mask = torch.full((S, S), torch.finfo(torch.float32).min).triu(1)
        s = s + mask
    return torch.softmax(s, dim=-1, dtype=torch.float32)


# --------------------------------------------------------------------------------------
# (a6) selection
# --------------------------------------------------------------------------------------
def num_selected(ratio: float, S: int) -> int:
    """k = int(ratio * S): Python float product, truncation.  pythia_model.py:170."""
    return int(ratio * S)


def least_important(imp: torch.Tensor, k: int) -> torch.Tensor:
    """argsort ascending, first k; ties resolved by lower index (north-star rule).
    pythia_model.py:169-170, qwen_layer_wise.py:57 (reference uses the default, unstable sort)."""
    return torch.argsort(imp, descending=False, stable=True)[:k]


def ladder_bits(imp: torch.Tensor, ladder: Sequence[Tuple[int, int]]) -> np.ndarray:
    """Bit-width map.  `ladder` = [(k0,b0),(k1,b1),...]: the k0 least-important tokens get b0
    bits, the next k1 get b1, ...; everything else crosses as raw fp32 (32)."""
    S = imp.numel()
    order = torch.argsort(imp, descending=False, stable=True).numpy()
    bits = np.full(S, BITS_RAW, dtype=np.uint8)
    pos = 0
    for k, b in ladder:
        bits[order[pos:pos + k]] = b
        pos += k
    return bits


# --------------------------------------------------------------------------------------
# (a7-a9) scale, quantize, dequantize -- reference op order, fp32
# --------------------------------------------------------------------------------------
def quant_codes(x: torch.Tensor, m: torch.Tensor, L: int) -> torch.Tensor:
    """round_half_even(clamp(x / m * L, -(L+1), L)) kept as fp32 integers, exactly the op
    sequence of pythia_model.py:180-184 / qwen_layer_wise.py:64-66 (L=7 there)."""
    scaled = torch.clamp(x / m * float(L), -float(L + 1), float(L))
    return torch.round(scaled)


def dequant_values(q: torch.Tensor, m: torch.Tensor, L: int, recip: bool = False) -> torch.Tensor:
    """q / L * m  (pythia_model.py:186, qwen_layer_wise.py:68).  recip=True restates what CUDA ATen
    computes for the same source line: a tensor divided by a Python scalar is multiplied by the fp32
    reciprocal of the scalar there (aten/src/ATen/native/cuda/BinaryDivTrueKernel.cu), so the value
    is (q * fl(1/L)) * m -- pinned against torch on a GPU by tests/test_cuda_aten_parity_gpu.py."""
    if recip:
        return q * torch.tensor(1.0, dtype=torch.float32).div(torch.tensor(float(L), dtype=torch.float32)) * m
    return q / float(L) * m


def fake_quant_reference(h: torch.Tensor, imp: torch.Tensor, ratio: float) -> torch.Tensor:
    """The reference's whole boundary block: int4, ONE absmax for the selected set, in place
    on (1,S,D).  pythia_model.py:167-188 == qwen_layer_wise.py:54-70.  Returns the new tensor."""
    h = h.clone()
    if not ratio > 0:
        return h
    idx = least_important(imp, num_selected(ratio, h.shape[1]))
    m = torch.max(torch.abs(h[:, idx, :]))
    q = quant_codes(h[:, idx, :], m, 7)
    h[:, idx, :] = dequant_values(q, m, 7)
    return h


def reference_boundary_cpu(A: torch.Tensor, h: torch.Tensor, ladder, scale_mode: int, method="regular_importance"):
    """The reference's boundary path as it runs on a CPU, vectorised ATen ops only: importance
    (last_row_exp.py:13-30), argsort selection (pythia_model.py:169-170) and, per ladder rung, the
    gather / absmax / scale / clamp / round / rescale / index_put block (pythia_model.py:176-188)
    applied in place.  Used as the timed CPU baseline; it never packs (the reference does not)."""
    imp = importance_regular(A) if method == "regular_importance" else importance_last_row(A)
    order = torch.argsort(imp, descending=False)
    pos = 0
    for k, b in ladder:
        idx = order[pos:pos + k]
        pos += k
        if k == 0:
            continue
        L = float(levels(b))
        sel = h[:, idx, :]
        m = torch.max(torch.abs(sel)) if scale_mode == SCALE_GROUP else sel.abs().amax(dim=-1, keepdim=True)
        q = torch.round(torch.clamp(sel / m * L, -(L + 1.0), L))
        h[:, idx, :] = q / L * m
    return imp, h


def reference_boundary_cpu_from_qk(q, k, h, scaling, ladder, scale_mode, method="regular_importance"):
    """Same as reference_boundary_cpu, starting where the fused GPU path starts: the split layer's
    post-RoPE Q and K.  The reference gets its attention map from HF eager attention
    (last_row_exp.py:92-95), which is exactly attention_probs()."""
    A = attention_probs(q, k, scaling, causal=True)
    return reference_boundary_cpu(A, h, ladder, scale_mode, method)


def channel_fake_quant_reference(h: torch.Tensor, method: str) -> torch.Tensor:
    """Channel-wise quantise->dequantise of (B,S,D), the arithmetic of qwen_layer_wise.py:125-152.
    absmax variants are vectorised over channels (max is order independent); the mean variant
    keeps the reference's per-channel torch.mean over the strided column so the scale is bit-equal."""
    h = h.clone()
    if method in ("channel_8", "channel_4"):
        L = 127 if method == "channel_8" else 7
        m = h.abs().amax(dim=(0, 1))
        q = torch.round(h / m * L)
        return q * m / L
    if method == "channel_1_max":
        m = h.abs().amax(dim=(0, 1))
        q = torch.clamp(torch.round(h / m), -1, 1)
        return q * m
    if method == "channel_1_mean":
        for c in range(h.shape[2]):
            col = h[:, :, c]
            u = torch.mean(col) + 1e-8
            h[:, :, c] = torch.clamp(torch.round(col / u), -1, 1) * u
        return h
    raise ValueError(f"Unknown method: {method}")


# --------------------------------------------------------------------------------------
# packed payload (our format; see header comment)
# --------------------------------------------------------------------------------------
def payload_layout(S: int, D: int, scale_mode: int) -> Tuple[int, int, int, int]:
    """(bits_off, rowoff_off, scales_off, codes_off) byte offsets inside one window's payload."""
    bits_off = HEADER_BYTES
    rowoff_off = bits_off + pad16(S)
    scales_off = rowoff_off + pad16(4 * S)
    n_scales = GROUP_SLOTS if scale_mode == SCALE_GROUP else S
    codes_off = scales_off + pad16(4 * n_scales)
    return bits_off, rowoff_off, scales_off, codes_off


def payload_bytes(S: int, D: int, counts: Sequence[Tuple[int, int]], scale_mode: int) -> int:
    """Total bytes; depends only on how many tokens take each width."""
    codes_off = payload_layout(S, D, scale_mode)[3]
    n = codes_off
    used = 0
    for k, b in counts:
        n += k * row_bytes(D, b)
        used += k
    n += (S - used) * row_bytes(D, BITS_RAW)
    return n


def _pack_row(codes: np.ndarray, bits: int) -> np.ndarray:
    """Two's-complement b-bit fields, element e at bit e*b, little-endian bit order."""
    D = codes.shape[0]
    field = (codes.astype(np.int64) & ((1 << bits) - 1)).astype(np.uint8)
    per = 8 // bits
    padded = np.zeros(((D + per - 1) // per) * per, dtype=np.uint8)
    padded[:D] = field
    padded = padded.reshape(-1, per)
    out = np.zeros(padded.shape[0], dtype=np.uint8)
    for i in range(per):
        out |= padded[:, i] << (bits * i)
    return out


def _unpack_row(raw: np.ndarray, bits: int, D: int) -> np.ndarray:
    per = 8 // bits
    mask = (1 << bits) - 1
    fields = np.stack([(raw >> (bits * i)) & mask for i in range(per)], axis=1).reshape(-1)[:D].astype(np.int16)
    sign = 1 << (bits - 1)
    return ((fields ^ sign) - sign).astype(np.int8)


def token_scales(h: torch.Tensor, bits: np.ndarray, scale_mode: int) -> np.ndarray:
    """Scale table as stored in the payload.  group: slot per width = absmax over all tokens of
    that width (reference a7 generalised per class); token: absmax of each row (raw rows: 0)."""
    S = h.shape[0]
    if scale_mode == SCALE_GROUP:
        out = np.zeros(GROUP_SLOTS, dtype=np.float32)
        for b in (8, 4, 2):
            sel = np.nonzero(bits == b)[0]
            if sel.size:
                out[slot_of_bits(b)] = torch.max(torch.abs(h[torch.from_numpy(sel)])).item()
        return out
    out = torch.max(torch.abs(h), dim=1).values.numpy().astype(np.float32).copy()
    out[bits == BITS_RAW] = 0.0
    return out


def row_codes(h_row: torch.Tensor, m: float, bits: int, scale_mode: int) -> np.ndarray:
    """int8 codes of one row.  NaN (0/0 when m == 0) cannot be stored: code 0."""
    L = levels(bits)
    q = quant_codes(h_row, torch.tensor(m, dtype=torch.float32), L)
    q = torch.nan_to_num(q, nan=0.0)
    return q.to(torch.int8).numpy()


SCALE_CHANNEL_BASE = 2  # header scale_mode of a per-channel payload = SCALE_CHANNEL_BASE + index in CHANNEL_METHODS
CHANNEL_METHODS = ("channel_8", "channel_4", "channel_1_mean", "channel_1_max")
CHANNEL_BITS = {"channel_8": 8, "channel_4": 4, "channel_1_mean": 2, "channel_1_max": 2}


def channel_codes_reference(h: torch.Tensor, method: str):
    """(per-channel statistic (D,), integer codes (S,D) as fp32) of ONE window (S,D): the first halves of the
    formulas of qwen_layer_wise.py:132,139-141,147-148 (channel_fake_quant_reference is codes -> floats)."""
    if method in ("channel_8", "channel_4"):
        L = 127 if method == "channel_8" else 7
        m = h.abs().amax(dim=0)
        return m, torch.round(h / m * L)
    if method == "channel_1_max":
        m = h.abs().amax(dim=0)
        return m, torch.clamp(torch.round(h / m), -1, 1)
    if method == "channel_1_mean":
        u = torch.stack([torch.mean(h[:, c]) + 1e-8 for c in range(h.shape[1])])
        return u, torch.clamp(torch.round(h / u), -1, 1)
    raise ValueError(f"Unknown method: {method}")


def pack_payload_channel(h: torch.Tensor, method: str) -> np.ndarray:
    """One window (S,D) -> per-channel payload: same sections as pack_payload, scales = D statistics, every
    row at CHANNEL_BITS[method] bits, rows back to back."""
    S, D = h.shape
    mode = SCALE_CHANNEL_BASE + CHANNEL_METHODS.index(method)
    b = CHANNEL_BITS[method]
    bits_off = HEADER_BYTES
    rowoff_off = bits_off + pad16(S)
    scales_off = rowoff_off + pad16(4 * S)
    codes_off = scales_off + pad16(4 * D)
    rb = row_bytes(D, b)
    total = codes_off + S * rb
    buf = np.zeros(total, dtype=np.uint8)
    buf[:HEADER_BYTES] = np.frombuffer(struct.pack("<8I", MAGIC, mode, S, D, D, codes_off, total, 0), dtype=np.uint8)
    buf[bits_off:bits_off + S] = b
    buf[rowoff_off:rowoff_off + 4 * S] = (np.arange(S, dtype=np.uint32) * (rb // 16)).view(np.uint8)
    stat, q = channel_codes_reference(h, method)
    buf[scales_off:scales_off + 4 * D] = stat.numpy().astype(np.float32).view(np.uint8)
    codes = torch.nan_to_num(q, nan=0.0).to(torch.int8).numpy()
    for t in range(S):
        packed = _pack_row(codes[t], b)
        buf[codes_off + t * rb:codes_off + t * rb + packed.size] = packed
    return buf


def unpack_payload_channel(buf: np.ndarray) -> torch.Tensor:
    """Dequantise a per-channel payload with the reference's operation order: q * m / L (absmax widths),
    q * u (ternary); a zero statistic gives NaN like the reference's division by it."""
    magic, mode, S, D, n_scales, codes_off, total, _ = struct.unpack("<8I", buf[:HEADER_BYTES].tobytes())
    assert magic == MAGIC and total == buf.size and n_scales == D
    method = CHANNEL_METHODS[mode - SCALE_CHANNEL_BASE]
    b = CHANNEL_BITS[method]
    scales_off = HEADER_BYTES + pad16(S) + pad16(4 * S)
    stat = torch.from_numpy(buf[scales_off:scales_off + 4 * D].view(np.float32).copy())
    rb = row_bytes(D, b)
    q = torch.stack([torch.from_numpy(_unpack_row(buf[codes_off + t * rb:codes_off + t * rb + (D * b + 7) // 8], b, D)
                                      .astype(np.float32)) for t in range(S)])
    out = q * stat / (127.0 if method == "channel_8" else 7.0) if method in ("channel_8", "channel_4") else q * stat
    out[:, stat == 0] = float("nan")
    return out


HDR_DEQUANT_RECIP = 1  # header word 7, bit 0: the receiver dequantises with q * fl(1/L) (CUDA ATen flavour)


def pack_payload(h: torch.Tensor, bits: np.ndarray, scale_mode: int, recip: bool = False) -> np.ndarray:
    """h: (S,D) fp32, bits: (S,) u8 in {2,4,8,32} -> uint8 payload for one window."""
    S, D = h.shape
    bits_off, rowoff_off, scales_off, codes_off = payload_layout(S, D, scale_mode)
    total = codes_off + int(sum(row_bytes(D, int(b)) for b in bits))
    buf = np.zeros(total, dtype=np.uint8)
    n_scales = GROUP_SLOTS if scale_mode == SCALE_GROUP else S
    buf[:HEADER_BYTES] = np.frombuffer(
        struct.pack("<8I", MAGIC, scale_mode, S, D, n_scales, codes_off, total, HDR_DEQUANT_RECIP if recip else 0),
        dtype=np.uint8)
    buf[bits_off:bits_off + S] = bits
    rb = np.array([row_bytes(D, int(b)) for b in bits], dtype=np.int64)
    row_off16 = ((np.cumsum(rb) - rb) // 16).astype(np.uint32)
    buf[rowoff_off:rowoff_off + 4 * S] = row_off16.view(np.uint8)
    scales = token_scales(h, bits, scale_mode)
    buf[scales_off:scales_off + 4 * n_scales] = scales.view(np.uint8)
    off = codes_off
    for t in range(S):
        b = int(bits[t])
        if b == BITS_RAW:
            raw = h[t].numpy().view(np.uint8)
            buf[off:off + raw.size] = raw
        else:
            m = float(scales[slot_of_bits(b)] if scale_mode == SCALE_GROUP else scales[t])
            packed = _pack_row(row_codes(h[t], m, b, scale_mode), b)
            buf[off:off + packed.size] = packed
        off += row_bytes(D, b)
    return buf


def unpack_payload(buf: np.ndarray) -> torch.Tensor:
    """Inverse of pack_payload: dequantised (S,D) fp32 hidden states.
    group mode with a zero class scale reproduces the reference's unguarded 0/0 -> NaN rows
    (pythia_model.py:176-186); token mode treats an all-zero row as zeros (our extension)."""
    magic, scale_mode, S, D, n_scales, codes_off, total, flags = struct.unpack("<8I", buf[:HEADER_BYTES].tobytes())
    assert magic == MAGIC and total == buf.size
    bits_off, rowoff_off, scales_off, _ = payload_layout(S, D, scale_mode)
    bits = buf[bits_off:bits_off + S]
    row_off16 = buf[rowoff_off:rowoff_off + 4 * S].view(np.uint32)
    scales = buf[scales_off:scales_off + 4 * n_scales].view(np.float32)
    out = torch.empty(S, D, dtype=torch.float32)
    for t in range(S):
        b = int(bits[t])
        off = codes_off + 16 * int(row_off16[t])
        if b == BITS_RAW:
            out[t] = torch.from_numpy(buf[off:off + 4 * D].view(np.float32).copy())
        else:
            L = levels(b)
            m = float(scales[slot_of_bits(b)] if scale_mode == SCALE_GROUP else scales[t])
            q = torch.from_numpy(_unpack_row(buf[off:off + (D * b + 7) // 8], b, D).astype(np.float32))
            if scale_mode == SCALE_GROUP and m == 0.0:
                out[t] = float("nan")
            else:
                out[t] = dequant_values(q, torch.tensor(m, dtype=torch.float32), L, recip=bool(flags & HDR_DEQUANT_RECIP))
    return out


def boundary_roundtrip(h: torch.Tensor, imp: torch.Tensor, ladder, scale_mode: int, recip: bool = False):
    """score-independent part of the path: select -> quantize/pack -> (transfer) -> dequantize."""
    bits = ladder_bits(imp, ladder)
    payload = pack_payload(h, bits, scale_mode, recip)
    return bits, payload, unpack_payload(payload)


# --------------------------------------------------------------------------------------
# downstream first norm (fused with dequant in K4)
# --------------------------------------------------------------------------------------
def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float) -> torch.Tensor:
    """nn.LayerNorm as used by GPT-NeoX layers (third-party HF; called through
    pythia_model.py:165 on the layer after the split)."""
    return torch.nn.functional.layer_norm(x, (x.shape[-1],), gamma, beta, eps)


def rms_norm(x: torch.Tensor, gamma: torch.Tensor, eps: float) -> torch.Tensor:
    """Qwen2RMSNorm (third-party HF; reached through qwen_layer_wise.py:52)."""
    var = x.pow(2).mean(-1, keepdim=True)
    return gamma * (x * torch.rsqrt(var + eps))


# --------------------------------------------------------------------------------------
# (8f-4) layer-distance analytics, numpy like the notebook
# --------------------------------------------------------------------------------------
def kl_divergence_np(p: np.ndarray, q: np.ndarray) -> float:
    """Notebooks/distributions_distance_across_layers.ipynb cell 12."""
    with np.errstate(divide="ignore", invalid="ignore"):
        return float(np.sum(np.where(p != 0, p * np.log2(p / q), 0)))


def jensen_shannon_np(p: np.ndarray, q: np.ndarray) -> float:
    """Same notebook, cell 13 (`jenson_shannon_distance`)."""
    m = 0.5 * (p + q)
    return 0.5 * (kl_divergence_np(p, m) + kl_divergence_np(q, m))


def layer_distance_matrix_np(dist: np.ndarray) -> np.ndarray:
    """Same notebook, cell 16: upper triangle of sample-averaged JS divergences."""
    L, N = dist.shape[0], dist.shape[1]
    out = np.full((L, L), np.nan)
    for i in range(L):
        for j in range(i + 1, L):
            out[i, j] = sum(jensen_shannon_np(dist[i][s], dist[j][s]) for s in range(N)) / N
    return out


MASS_ONE = 1 << 47  # fixed-point unit of the top-rho mass sums


def mass_threshold_fixed(threshold: float) -> int:
    """ceil(threshold * 2^47) as an integer (0 for thresholds <= 0)."""
    if threshold <= 0:
        return 0
    scaled = float(threshold) * float(MASS_ONE)
    q = int(scaled)
    return q + 1 if q < scaled else q


def top_rho_keep_count_fixed(distribution: np.ndarray, threshold: float) -> int:
    """The intended rule of batched_top_rho_quantization (pythia_model.py:93-106) -- walk the tokens
    from the most important down and take one while the mass taken so far is below the threshold
    (the reference's threshold is 1 - 0.1*ratio) -- in the exact arithmetic the kernel uses: each
    score, clamped to [0, 1], counts as the integer floor(score * 2^47), so the result does not
    depend on the order of summation."""
    v = np.asarray(distribution, dtype=np.float32)
    mass = np.floor(np.clip(np.nan_to_num(v, nan=0.0), 0.0, 1.0).astype(np.float64) * float(MASS_ONE)).astype(np.uint64)
    order = np.argsort(-v.astype(np.float64), kind="stable")
    T, total, n = mass_threshold_fixed(threshold), 0, 0
    for i in order:
        if total >= T:
            break
        total += int(mass[i])
        n += 1
    return n


def top_rho_keep_count(distribution: np.ndarray, ratio: float) -> int:
    """pythia_model.py:93-106: iterate the descending distribution until the running total reaches
    1 - 0.1*ratio; returns how many tokens were taken."""
    threshold = 1 - 0.1 * ratio
    total, n = 0, 0
    for value in sorted(distribution.tolist(), reverse=True):
        if total >= threshold:
            break
        total += value
        n += 1
    return n


# --------------------------------------------------------------------------------------
# (a12) sliding-window schedule and accounting
# --------------------------------------------------------------------------------------
def window_schedule(n_tokens: int, max_length: int, stride: int):
    """[(begin, end, trg_len)] exactly as the loop at last_row_exp.py:85-90,115,129-130."""
    out = []
    prev_end = 0
    for begin in range(0, n_tokens, stride):
        end = min(begin + max_length, n_tokens)
        out.append((begin, end, end - prev_end))
        prev_end = end
        if end == n_tokens:
            break
    return out


def window_targets(ids: torch.Tensor, trg_len: int) -> Tuple[torch.Tensor, int]:
    """labels with -100 on context, and num_loss_tokens (last_row_exp.py:89-90,100-102)."""
    tgt = ids.clone()
    tgt[:, :-trg_len] = -100
    valid = int((tgt != -100).sum().item())
    return tgt, valid - tgt.size(0)


def shifted_nll(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """pythia_model.py:197-203, qwen_layer_wise.py:32-37."""
    t = target[:, 1:]
    lg = logits[:, :-1, :]
    return torch.nn.functional.cross_entropy(lg.reshape(-1, lg.size(-1)), t.reshape(-1), ignore_index=-100)


def perplexity(total_nll: float, n_tokens: int) -> float:
    """last_row_exp.py:138."""
    return math.exp(total_nll / n_tokens)


# --------------------------------------------------------------------------------------
# whole-window reference sweep on the CPU (timed baseline for the eval driver)
# --------------------------------------------------------------------------------------
def reference_window_sweep_cpu(model, eager_model, ids, target, methods, layers_of_interest, ratios):
    """What one iteration of the reference loop does for one window (last_row_exp.py:92-111):
    an eager forward for the attention maps, get_importance_order per method, then a FULL
    layer-by-layer forward with the in-place int4 block for every (method, layer, ratio)
    (pythia_model.py:153-206 / qwen_layer_wise.py:42-76).  HF modules on the CPU; returns the
    (M, L, R) NLL tensor."""
    base = model.base_model
    embed = getattr(base, "embed_in", None) or base.embed_tokens
    final_norm = getattr(base, "final_layer_norm", None) or base.norm
    head = getattr(model, "embed_out", None) or model.lm_head
    num_layers = len(base.layers)
    out = torch.zeros(len(methods), len(layers_of_interest), len(ratios))
    with torch.no_grad():
        att = eager_model(input_ids=ids, output_attentions=True).attentions
        imps = {m: get_importance_order(m, att, num_layers) for m in methods}
        for mi, m in enumerate(methods):
            for li, loi in enumerate(layers_of_interest):
                for ri, ratio in enumerate(ratios):
                    h = embed(ids)
                    pos_ids = torch.arange(ids.size(1)).unsqueeze(0)
                    pos = base.rotary_emb(h, pos_ids)
                    for i in range(num_layers):
                        o = base.layers[i](h, position_embeddings=pos)
                        h = o[0] if isinstance(o, (tuple, list)) else o
                        if i == loi and ratio > 0:
                            h = fake_quant_reference(h, imps[m][loi], ratio)
                    out[mi, li, ri] = shifted_nll(head(final_norm(h)), target)
    return out
