"""GPT-2 small (north-star config ``BASELINE.json:configs[4]``: DDP bf16, seq 1024, large gradient
buckets fused with the optimizer step; not part of the reference).  124,439,808 parameters with the
LM head tied to the token embedding.  Linear layers run the tcgen05 GEMM; attention is the tcgen05 flash-attention kernel
of csrc/ops/attention_tc.cu for bf16 / head dim 64 / T % 128 == 0 (the library SDPA call otherwise, and in the stock arm)."""
from __future__ import annotations

import math
import os
from dataclasses import dataclass

import torch
import torch.nn.functional as F
from torch import nn

from .. import ops


def cuda_ops():
    from ..ops import cuda
    return cuda


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12

    @property
    def padded_vocab(self) -> int:
        """Vocabulary rounded up to a multiple of 128 so the LM-head GEMMs tile on tensor cores; the extra logits are
        never targets (124,439,808 parameters are counted on the unpadded 50257 x 768 embedding)."""
        return -(-self.vocab_size // 128) * 128


class _SplitHeads(torch.autograd.Function):
    """qkv [B*T, 3C] -> q, k, v as [B, H, T, D] strided views (no copies forward).  Backward writes dq / dk / dv straight into one
    [B, T, 3, H, D] buffer (three strided copies) instead of autograd's zero-fill + add per selected slice."""

    @staticmethod
    def forward(ctx, qkv, B, T, H):
        D = qkv.shape[1] // (3 * H)
        ctx.dims = (B, T, H, D)
        v = qkv.view(B, T, 3, H, D)
        return v[:, :, 0].transpose(1, 2), v[:, :, 1].transpose(1, 2), v[:, :, 2].transpose(1, 2)

    @staticmethod
    def backward(ctx, dq, dk, dv):
        B, T, H, D = ctx.dims
        out = torch.empty(B, T, 3, H, D, dtype=dq.dtype, device=dq.device)
        out[:, :, 0].copy_(dq.transpose(1, 2))
        out[:, :, 1].copy_(dk.transpose(1, 2))
        out[:, :, 2].copy_(dv.transpose(1, 2))
        return out.view(B * T, 3 * H * D), None, None, None


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.ln_1 = ops.LayerNorm(cfg.n_embd)
        self.attn_qkv = ops.Linear(cfg.n_embd, 3 * cfg.n_embd)
        self.attn_proj = ops.Linear(cfg.n_embd, cfg.n_embd)
        self.ln_2 = ops.LayerNorm(cfg.n_embd)
        self.act = ops.GELU()
        self.mlp_fc = ops.Linear(cfg.n_embd, 4 * cfg.n_embd)
        self.mlp_proj = ops.Linear(4 * cfg.n_embd, cfg.n_embd)
        self.n_head = cfg.n_head

    def _attention(self, h: torch.Tensor, B: int, T: int, C: int) -> torch.Tensor:
        qkv = self.attn_qkv(h.view(B * T, C))
        if ops.layers.use_kernels(qkv) and os.environ.get("DCP_ATTENTION", "1") == "1" and cuda_ops().causal_attention_supported(qkv, T, self.n_head):
            # flash attention on tcgen05, reading q / k / v in place and returning the heads merged (csrc/ops/attention_tc.cu)
            y = cuda_ops().causal_attention(qkv, B, T, self.n_head)
        else:
            q, k, v = _SplitHeads.apply(qkv, B, T, self.n_head)
            y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
            y = y.transpose(1, 2).reshape(B * T, C)
        return self.attn_proj(y).view(B, T, C)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, T, C = x.shape
        x = x + self._attention(self.ln_1(x), B, T, C)
        h = self.act(self.mlp_fc(self.ln_2(x).view(B * T, C)))
        return x + self.mlp_proj(h).view(B, T, C)

    def forward_fused(self, x: torch.Tensor, pending):
        """Same block with every residual add folded into the LayerNorm that follows it (ops.cuda.add_layer_norm): takes the
        residual stream ``x`` and the previous block's MLP output ``pending`` that has not been added yet (None for the first
        block), returns the stream after the attention add and this block's own MLP output, again un-added."""
        K = cuda_ops()
        B, T, C = x.shape
        if pending is None:
            h = self.ln_1(x)
        else:
            x, h = K.add_layer_norm(x, pending, self.ln_1.weight, self.ln_1.bias, self.ln_1.eps)
        a = self._attention(h, B, T, C)
        x, h = K.add_layer_norm(x, a, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
        return x, self.mlp_proj(self.act(self.mlp_fc(h.view(B * T, C)))).view(B, T, C)


class GPT2(nn.Module):
    def __init__(self, cfg: GPT2Config = GPT2Config()):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.padded_vocab, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList(Block(cfg) for _ in range(cfg.n_layer))
        self.ln_f = ops.LayerNorm(cfg.n_embd)
        self.compute_dtype = None   # torch.bfloat16: bf16 activations / GEMMs with fp32 master weights
        self.apply(self._init)
        for n, p in self.named_parameters():
            if n.endswith("proj.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward(self, idx: torch.Tensor, targets: torch.Tensor = None):
        B, T = idx.shape
        kern = ops.layers.use_kernels(idx) and os.environ.get("DCP_GPT2_FUSED", "1") == "1"
        if kern and self.wte.weight.dtype == torch.float32:
            # one gather kernel: wte[idx] + wpe[pos], written in the compute dtype (csrc/ops/transformer.cu)
            x = cuda_ops().embed(idx, self.wte.weight, self.wpe.weight, self.compute_dtype or torch.float32)
        else:
            pos = torch.arange(T, device=idx.device)
            x = self.wte(idx) + self.wpe(pos)
            if self.compute_dtype is not None:
                x = x.to(self.compute_dtype)
        if kern and cuda_ops().add_layer_norm_supported(x):
            pending = None
            for blk in self.h:
                x, pending = blk.forward_fused(x, pending)
            _, x = cuda_ops().add_layer_norm(x, pending, self.ln_f.weight, self.ln_f.bias, self.ln_f.eps)
        else:
            for blk in self.h:
                x = blk(x)
            x = self.ln_f(x)
        if ops.layers.use_kernels(x):
            from ..ops import cuda
            logits = cuda.linear(x.view(B * T, -1), self.wte.weight, None, False)
        else:
            logits = F.linear(x.view(B * T, -1), self.wte.weight)
        if targets is None:
            return logits.view(B, T, -1)
        return ops.cross_entropy(logits.float() if not logits.is_cuda else logits, targets.reshape(-1))
