"""Causal flash attention on tcgen05 / TMEM (csrc/ops/attention_tc.cu) against a plain PyTorch fp32 attention of the same
inputs: forward output, log-sum-exp, and all three gradients, on shapes that cover a single tile, several key blocks per query
block, several heads / batches, and the GPT-2 small shape (12 heads, T = 1024); then the GPT-2 block with the kernel on and off."""
import copy
import os

import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200.ops import cuda as K

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _reference(qkv, dout, B, T, H):
    C = H * 64
    x = qkv.float().clone().requires_grad_(True)
    v = x.view(B, T, 3, H, 64)
    q, k, vv = v[:, :, 0].transpose(1, 2), v[:, :, 1].transpose(1, 2), v[:, :, 2].transpose(1, 2)
    s = (q @ k.transpose(-1, -2)) * 0.125
    mask = torch.ones(T, T, device=qkv.device, dtype=torch.bool).tril()
    p = torch.softmax(s.masked_fill(~mask, float("-inf")), -1)
    out = (p @ vv).transpose(1, 2).reshape(B * T, C)
    out.backward(dout.float())
    return out.detach(), x.grad


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (1, 384, 2), (2, 1024, 3), (4, 1024, 12)])
def test_flash_attention_forward_backward_match_fp32(B, T, H):
    C = H * 64
    g = torch.Generator(device=DEV).manual_seed(B * T + H)
    qkv = torch.randn(B * T, 3 * C, device=DEV, generator=g).bfloat16().requires_grad_(True)
    dout = torch.randn(B * T, C, device=DEV, generator=g).bfloat16()
    assert K.causal_attention_supported(qkv, T, H)
    out = K.causal_attention(qkv, B, T, H)
    out.backward(dout)
    torch.cuda.synchronize()
    ref_out, ref_grad = _reference(qkv.detach(), dout, B, T, H)
    assert not torch.isnan(out).any() and not torch.isnan(qkv.grad).any()
    assert (out.float() - ref_out).abs().max().item() < 3e-2                      # bf16 output of O(1) values
    for name, sl in (("dq", slice(0, C)), ("dk", slice(C, 2 * C)), ("dv", slice(2 * C, 3 * C))):
        a, b = qkv.grad[:, sl].float(), ref_grad[:, sl]
        assert ((a - b).norm() / b.norm()).item() < 1e-2, name                    # bf16 P / dS operands: ~0.3 % in norm


def test_gpt2_block_with_and_without_the_attention_kernel():
    torch.manual_seed(0)
    cfg = dcp.models.GPT2Config(vocab_size=512, n_positions=256, n_embd=128, n_layer=2, n_head=2)     # head dim 64
    m = dcp.models.GPT2(cfg).to(DEV)
    m.compute_dtype = torch.bfloat16
    m2 = copy.deepcopy(m)
    idx = torch.randint(0, 512, (2, 256), device=DEV); tgt = torch.randint(0, 512, (2, 256), device=DEV)
    before = K.launch_count
    la = m(idx, tgt); la.backward()
    os.environ["DCP_ATTENTION"] = "0"
    try:
        lb = m2(idx, tgt); lb.backward()
    finally:
        os.environ.pop("DCP_ATTENTION")
    assert K.launch_count > before
    assert abs(la.item() - lb.item()) < 2e-2 * abs(lb.item())
    for (n, p), q in zip(m.named_parameters(), m2.parameters()):
        rel = ((p.grad - q.grad).norm() / (q.grad.norm() + 1e-12)).item()
        assert rel < 0.08, (n, rel)                                                # two bf16 attention implementations, two layers deep
