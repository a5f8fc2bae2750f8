"""Vision Transformer with the reference's module API, running on hand-written sm_100a kernels.

Drop-in for the reference `models/vit.py`: same five classes, same constructor arguments (names, order,
defaults), same `forward(x)` signatures, same sub-module attribute names -- hence the same `state_dict()`
keys and shapes (152 fp32 tensors for ViT-B/16) and the same random-number consumption at construction, so
`torch.manual_seed(s); ViT(...)` yields bit-identical initial weights. The parameter containers are the very
torch modules the reference instantiates (nn.Conv2d, nn.LayerNorm, nn.MultiheadAttention, nn.Linear); only
their arithmetic is replaced: no forward of those containers is ever called, every `forward` below goes to
libb200vit.so through b200vit.functional.

Semantics kept from the reference (file:line under the reference repo):
  PatchEmbedding              models/vit.py:5-67     conv k=s=patch as GEMM, class token, + position, dropout
  MultiHeadSelfAttentionBlock models/vit.py:69-98    LN -> MHA(q=k=v), returns the branch WITHOUT the residual
  MLPBlock                    models/vit.py:100-131  LN -> Linear -> GELU(erf) -> Dropout -> Linear -> Dropout
  TransformerEncoderBlock     models/vit.py:133-169  x = msa(x) + x ; x = mlp(x) + x
  ViT                         models/vit.py:172-236  embed -> L blocks -> LN -> Linear on token 0
"""
from __future__ import annotations

import torch
from torch import nn

from .. import fp32 as FP_
from .. import functional as F_


def _check_geometry(embedding_dim: int, num_heads: int) -> None:
    """The attention kernels are written for 64-wide heads and the LayerNorm backward for rows of at most 1024
    values (ViT-B/16: 768 = 12 x 64, ViT-L/16: 1024 = 16 x 64). The reference accepts any pair; say so at construction
    instead of failing inside the first forward."""
    if embedding_dim != num_heads * F_.HEAD_DIM:
        raise NotImplementedError(
            f"b200vit supports head_dim == {F_.HEAD_DIM} only: embedding_dim {embedding_dim} with {num_heads} heads "
            f"gives {embedding_dim / num_heads:g}")
    if embedding_dim > F_.MAX_LN_DIM:
        raise NotImplementedError(f"b200vit supports embedding_dim <= {F_.MAX_LN_DIM} (got {embedding_dim})")


class PatchEmbedding(nn.Module):
    """Turns [B, C, H, W] images into the embedded token sequence [B, N+1, D]."""

    def __init__(self,
                 image_size: int,
                 color_channels: int = 3,
                 patch_size: int = 16,
                 embedding_dropout: float = 0.1,
                 embedding_dim: int = 768):
        super().__init__()
        self.patch_size = patch_size
        assert image_size % self.patch_size == 0, \
            f"Input image size must be divisible by patch size, image size: {image_size}, patch_size: {patch_size}"
        self.number_of_patches = int((image_size ** 2) / (patch_size ** 2))
        # Parameter containers in the reference's order (RNG: conv weight, conv bias, class token, positions).
        self.patch_and_flatten = nn.Sequential(
            nn.Conv2d(in_channels=color_channels, out_channels=embedding_dim, kernel_size=patch_size,
                      stride=patch_size),
            nn.Flatten(start_dim=2, end_dim=3),
        )
        self.class_token = nn.Parameter(torch.rand(1, 1, embedding_dim), requires_grad=True)
        self.position_embedding = nn.Parameter(torch.rand(1, self.number_of_patches + 1, embedding_dim),
                                               requires_grad=True)
        self.dropout = nn.Dropout(p=embedding_dropout)

    def forward(self, x: torch.Tensor):
        conv = self.patch_and_flatten[0]
        if F_.get_precision() == "fp32":
            return FP_.PatchEmbedF32Fn.apply(x, conv.weight, conv.bias, self.class_token, self.position_embedding,
                                             self.patch_size, self.dropout.p, self.training)
        return F_.PatchEmbedFn.apply(x, conv.weight, conv.bias, self.class_token, self.position_embedding,
                                     F_.bf16_shadow(conv.weight), self.patch_size, self.dropout.p, self.training,
                                     torch.is_grad_enabled())


class MultiHeadSelfAttentionBlock(nn.Module):
    """LayerNorm followed by multi-head self-attention (the residual is added by the caller)."""

    def __init__(self,
                 embedding_dim: int = 768,
                 num_heads: int = 12,
                 attn_dropout: float = 0):
        super().__init__()
        _check_geometry(embedding_dim, num_heads)
        self.layer_norm = nn.LayerNorm(normalized_shape=embedding_dim)
        self.multi_head_attention = nn.MultiheadAttention(embed_dim=embedding_dim, num_heads=num_heads,
                                                          dropout=attn_dropout, batch_first=True)

    def forward(self, x: torch.Tensor):
        mha = self.multi_head_attention
        if F_.get_precision() == "fp32":
            return FP_.MSAF32Fn.apply(x, self.layer_norm.weight, self.layer_norm.bias, mha.in_proj_weight,
                                      mha.in_proj_bias, mha.out_proj.weight, mha.out_proj.bias, mha.num_heads,
                                      mha.dropout, self.training, self.layer_norm.eps)
        # mha.dropout > 0 (attn_dropout): dropout on the attention probabilities in training mode, as torch does
        return F_.MSAFn.apply(x, self.layer_norm.weight, self.layer_norm.bias, mha.in_proj_weight, mha.in_proj_bias,
                              mha.out_proj.weight, mha.out_proj.bias, F_.bf16_shadow(mha.in_proj_weight),
                              F_.bf16_shadow(mha.out_proj.weight), mha.num_heads, mha.dropout, self.training,
                              self.layer_norm.eps)


class MLPBlock(nn.Module):
    """LayerNorm followed by the two-layer GELU MLP (the residual is added by the caller)."""

    def __init__(self,
                 embedding_dim: int = 768,
                 mlp_size=3072,
                 dropout: float = 0.1):
        super().__init__()
        self.layer_norm = nn.LayerNorm(normalized_shape=embedding_dim)
        self.mlp = nn.Sequential(
            nn.Linear(in_features=embedding_dim, out_features=mlp_size),
            nn.GELU(),
            nn.Dropout(p=dropout),
            nn.Linear(in_features=mlp_size, out_features=embedding_dim),
            nn.Dropout(p=dropout),
        )

    def forward(self, x: torch.Tensor):
        fc1, fc2 = self.mlp[0], self.mlp[3]
        if F_.get_precision() == "fp32":
            return FP_.MLPF32Fn.apply(x, self.layer_norm.weight, self.layer_norm.bias, fc1.weight, fc1.bias, fc2.weight,
                                      fc2.bias, self.mlp[2].p, self.training, self.layer_norm.eps)
        return F_.MLPFn.apply(x, self.layer_norm.weight, self.layer_norm.bias, fc1.weight, fc1.bias, fc2.weight,
                              fc2.bias, F_.bf16_shadow(fc1.weight), F_.bf16_shadow(fc2.weight), self.mlp[2].p,
                              self.training, torch.is_grad_enabled(), self.layer_norm.eps)


class TransformerEncoderBlock(nn.Module):
    """Pre-norm encoder block; runs as ONE fused autograd node (both residual adds live in GEMM epilogues)."""

    def __init__(self,
                 embedding_dim: int = 768,
                 num_heads: int = 12,
                 attn_dropout: float = 0,
                 mlp_size=3072,
                 mlp_dropout: float = 0.1):
        super().__init__()
        self.msa_block = MultiHeadSelfAttentionBlock(embedding_dim=embedding_dim, num_heads=num_heads,
                                                     attn_dropout=attn_dropout)
        self.mlp_block = MLPBlock(embedding_dim=embedding_dim, mlp_size=mlp_size, dropout=mlp_dropout)

    def forward(self, x: torch.Tensor):
        msa, mlp = self.msa_block, self.mlp_block
        mha = msa.multi_head_attention
        fc1, fc2 = mlp.mlp[0], mlp.mlp[3]
        if F_.get_precision() == "fp32":
            return FP_.EncoderBlockF32Fn.apply(
                x, msa.layer_norm.weight, msa.layer_norm.bias, mha.in_proj_weight, mha.in_proj_bias,
                mha.out_proj.weight, mha.out_proj.bias, mlp.layer_norm.weight, mlp.layer_norm.bias, fc1.weight,
                fc1.bias, fc2.weight, fc2.bias, mha.num_heads, mlp.mlp[2].p, self.training, mha.dropout,
                msa.layer_norm.eps, mlp.layer_norm.eps)
        return F_.EncoderBlockFn.apply(
            x, msa.layer_norm.weight, msa.layer_norm.bias, mha.in_proj_weight, mha.in_proj_bias,
            mha.out_proj.weight, mha.out_proj.bias, mlp.layer_norm.weight, mlp.layer_norm.bias, fc1.weight, fc1.bias,
            fc2.weight, fc2.bias, F_.bf16_shadow(mha.in_proj_weight), F_.bf16_shadow(mha.out_proj.weight),
            F_.bf16_shadow(fc1.weight), F_.bf16_shadow(fc2.weight), mha.num_heads, mlp.mlp[2].p, self.training,
            mha.dropout, torch.is_grad_enabled(), msa.layer_norm.eps, mlp.layer_norm.eps)


class ViT(nn.Module):
    """ViT classifier; defaults are ViT-Base/16 on 224x224 with a 1000-way head."""

    def __init__(self,
                 image_size: int = 224,
                 patch_size: int = 16,
                 num_transformer_layer: int = 12,
                 num_heads: int = 12,
                 embedding_dim: int = 768,
                 mlp_size: int = 3072,
                 attn_dropout: float = 0,
                 mlp_dropout: float = 0.1,
                 embedding_dropout: int = 0.1,
                 num_classes: int = 1000):
        super().__init__()
        # NB: like the reference, color_channels is not forwarded (3-channel images only through ViT).
        self.patch_embedding_block = PatchEmbedding(image_size=image_size, patch_size=patch_size,
                                                    embedding_dim=embedding_dim, embedding_dropout=embedding_dropout)
        self.transformer_encoder = nn.Sequential(*[
            TransformerEncoderBlock(embedding_dim=embedding_dim, num_heads=num_heads, attn_dropout=attn_dropout,
                                    mlp_size=mlp_size, mlp_dropout=mlp_dropout)
            for _ in range(num_transformer_layer)])
        self.layer_norm = nn.LayerNorm(normalized_shape=embedding_dim)
        self.classifier = nn.Sequential(nn.Linear(in_features=embedding_dim, out_features=num_classes))

    def forward(self, x: torch.Tensor):
        x = self.patch_embedding_block(x)
        x = self.transformer_encoder(x)
        head = self.classifier[0]
        return F_.HeadFn.apply(x, self.layer_norm.weight, self.layer_norm.bias, head.weight, head.bias,
                               self.layer_norm.eps)
