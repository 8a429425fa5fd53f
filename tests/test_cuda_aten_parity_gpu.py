"""The reference never runs on a CPU: every entry point hard-codes device = "cuda"
(Experiments/Pythia-70M/last_row_exp.py:57, initial_exp.py:78, Qwen2-0.5B/channel_wise.py:18).  The
oracle and the golden fixtures are CPU ATen.  This file runs the reference's fake-quant op sequence
(pythia_model.py:176-188 == qwen_layer_wise.py:60-70) with torch ON THE GPU and pins how the kernels
relate to it:

  * codes (`round(clamp(x / max_val * 7, -8, 7))`): bit-identical -- `x / max_val` divides by a
    0-dim CUDA tensor, which is a true division on both devices;
  * dequantised values (`q / 7 * max_val`): dividing by the Python scalar 7.0 is a multiplication
    by fl(1/7) in CUDA ATen, so the default (correctly rounded quotient, == CPU ATen) may differ by
    one ulp on some codes; a payload packed with cuda_aten_dequant=True is bit-identical.
"""
import numpy as np
import pytest
import torch

import splitq
import splitq_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def reference_block_on_cuda(h, idx):
    """The reference's statements, same order, same Python scalars, on CUDA tensors."""
    num_levels = 16
    max_val = torch.max(torch.abs(h[:, idx, :]))
    scaled_values = torch.clamp(h[:, idx, :] / max_val * (num_levels / 2 - 1), -(num_levels / 2), (num_levels / 2 - 1))
    quantized_values = torch.round(scaled_values)
    dequantized = quantized_values / (num_levels / 2 - 1) * max_val
    out = h.clone()
    out[:, idx, :] = dequantized
    return quantized_values, out, max_val


@pytest.mark.parametrize("S,D,ratio", [(512, 512, 0.5), (2048, 512, 0.25), (512, 896, 0.75), (333, 100, 1.0)])
def test_group_int4_block_vs_torch_on_cuda(S, D, ratio):
    gen = torch.Generator(device=DEV).manual_seed(S + D)
    h = torch.randn(1, S, D, device=DEV, generator=gen)
    h[:, :, :: max(1, D // 9)] *= 17.0
    imp = torch.rand(S, device=DEV, generator=gen)
    k = int(ratio * S)
    idx = torch.argsort(imp, descending=False, stable=True)[:k]
    q_ref, post_ref, max_val = reference_block_on_cuda(h, idx)

    ladder = splitq.ladder_for_ratio(ratio, S)
    nbytes = splitq.payload_bytes(S, D, ladder, O.SCALE_GROUP)
    bits, row_off = splitq.select_bits(imp, ladder, D)
    assert torch.equal((bits[0] == 4).nonzero().flatten(), torch.sort(idx).values)

    # codes: unpack the 4-bit fields of the selected rows and compare with torch's rounded values
    payload = splitq.quant_pack(h, bits, row_off, O.SCALE_GROUP, nbytes)
    buf = payload[0].cpu().numpy()
    codes_off = O.payload_layout(S, D, O.SCALE_GROUP)[3]
    offs = buf[O.payload_layout(S, D, O.SCALE_GROUP)[1]:][: 4 * S].view(np.uint32)
    rows = torch.sort(idx).values.cpu().numpy()
    q_sorted = q_ref[0][torch.argsort(idx)].cpu().numpy()  # reference rows in ascending token order
    for r, t in enumerate(rows):
        start = codes_off + 16 * int(offs[t])
        got = O._unpack_row(buf[start:start + (D * 4 + 7) // 8], 4, D)
        assert np.array_equal(got.astype(np.float32), q_sorted[r]), f"codes differ at token {t}"
    scale = buf[O.payload_layout(S, D, O.SCALE_GROUP)[2]:][:16].view(np.float32)[O.slot_of_bits(4)]
    assert scale == max_val.item()

    # dequantised values, CUDA-ATen flavour: bit-identical to torch on the GPU (up to the sign of zero:
    # an integer code has no -0)
    payload_r = splitq.quant_pack(h, bits, row_off, O.SCALE_GROUP, nbytes, cuda_aten_dequant=True)
    post_r = splitq.dequant(payload_r, S, D)
    assert torch.equal(post_r, post_ref)
    nz = post_ref != 0
    assert torch.equal(post_r.view(torch.int32)[nz], post_ref.view(torch.int32)[nz])
    # default flavour (== CPU ATen, correctly rounded q/7): equal or one ulp away, never more
    post_d = splitq.dequant(payload, S, D)
    ulp = (post_d.view(torch.int32) - post_ref.view(torch.int32)).abs()[nz]
    assert int(ulp.max()) <= 1
    frac = float((ulp != 0).float().mean())
    print(f"S={S} D={D} ratio={ratio}: default dequant differs from CUDA ATen by 1 ulp on {100 * frac:.2f}% of values")
    # and the default equals the true-division sequence evaluated on the same GPU
    seven = torch.tensor(7.0, device=DEV)
    true_div = (q_ref / seven * max_val)[0][torch.argsort(idx)]
    got_rows = post_d[0][torch.sort(idx).values]
    nzr = true_div != 0
    assert torch.equal(got_rows.view(torch.int32)[nzr], true_div.view(torch.int32)[nzr])
    # the oracle restates both flavours
    want_r = O.unpack_payload(O.pack_payload(h[0].cpu(), bits[0].cpu().numpy(), O.SCALE_GROUP, recip=True))
    assert np.array_equal(post_r[0].cpu().numpy().view(np.uint32), want_r.numpy().view(np.uint32))
    assert np.array_equal(payload_r[0].cpu().numpy(), O.pack_payload(h[0].cpu(), bits[0].cpu().numpy(), O.SCALE_GROUP, recip=True))


def test_reciprocal_flavour_covers_every_width_and_token_mode():
    S, D = 96, 256
    gen = torch.Generator(device=DEV).manual_seed(5)
    h = torch.randn(2, S, D, device=DEV, generator=gen) * 3
    imp = torch.rand(2, S, device=DEV, generator=gen)
    ladder = [(24, 2), (24, 4), (24, 8)]
    for mode in (O.SCALE_GROUP, O.SCALE_TOKEN):
        nbytes = splitq.payload_bytes(S, D, ladder, mode)
        bits, row_off = splitq.select_bits(imp, ladder, D)
        payload = splitq.quant_pack(h, bits, row_off, mode, nbytes, cuda_aten_dequant=True)
        out = splitq.dequant(payload, S, D).cpu()
        for b in range(2):
            want = O.unpack_payload(O.pack_payload(h[b].cpu(), bits[b].cpu().numpy(), mode, recip=True))
            assert np.array_equal(out[b].numpy().view(np.uint32), want.numpy().view(np.uint32)), (mode, b)
