"""splitq: B200-native split-inference boundary path (importance -> quantise/pack -> transfer ->
dequantise) behind the call surface of the reference's layer-wise wrappers."""
from . import _native
from ._native import (BITS_RAW, NORM_LAYERNORM, NORM_RMSNORM, SCALE_GROUP, SCALE_TOKEN, SplitqError)
from .boundary import (BoundaryPlan, HostPipeline, OverlappedBoundary, channel_fake_quant, channel_payload_bytes, channel_quant_pack, dequant, dequant_norm, importance_from_attention, importance_fused, ladder_for_ratio,
                       ladder_from_fractions, payload_bytes, quant_pack, quant_pack_group_noalloc, select_bits, transfer)
from .importance import get_importance_order

__all__ = [
    "BITS_RAW", "NORM_LAYERNORM", "NORM_RMSNORM", "SCALE_GROUP", "SCALE_TOKEN", "SplitqError", "BoundaryPlan", "HostPipeline", "OverlappedBoundary", "channel_fake_quant", "channel_payload_bytes", "channel_quant_pack",
    "dequant", "dequant_norm", "importance_from_attention", "importance_fused", "ladder_for_ratio", "ladder_from_fractions",
    "payload_bytes", "quant_pack", "quant_pack_group_noalloc", "select_bits", "transfer", "get_importance_order",
]
