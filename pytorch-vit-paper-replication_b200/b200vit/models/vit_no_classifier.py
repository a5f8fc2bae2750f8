"""Headless Vision Transformer: drop-in for the reference `models/vit_no_classifier.py`.

Same classes and constructor arguments as `b200vit.models.vit`, except that `ViT` has no `num_classes`, no
`classifier`, and `forward` returns the layer-normalised token sequence [B, N+1, D]
(reference models/vit_no_classifier.py:172-226). Used for feature extraction / forward-only inference.
"""
from __future__ import annotations

import torch
from torch import nn

from .. import functional as F_
from .vit import MLPBlock, MultiHeadSelfAttentionBlock, PatchEmbedding, TransformerEncoderBlock  # noqa: F401


class ViT(nn.Module):
    def __init__(self,
                 image_size: int = 224,
                 patch_size: int = 16,
                 num_transformer_layer: int = 12,
                 num_heads: int = 12,
                 embedding_dim: int = 768,
                 mlp_size: int = 3072,
                 attn_dropout: float = 0,
                 mlp_dropout: float = 0.1,
                 embedding_dropout: int = 0.1):
        super().__init__()
        self.patch_embedding_block = PatchEmbedding(image_size=image_size, patch_size=patch_size,
                                                    embedding_dim=embedding_dim, embedding_dropout=embedding_dropout)
        self.transformer_encoder = nn.Sequential(*[
            TransformerEncoderBlock(embedding_dim=embedding_dim, num_heads=num_heads, attn_dropout=attn_dropout,
                                    mlp_size=mlp_size, mlp_dropout=mlp_dropout)
            for _ in range(num_transformer_layer)])
        self.layer_norm = nn.LayerNorm(normalized_shape=embedding_dim)

    def forward(self, x: torch.Tensor):
        x = self.patch_embedding_block(x)
        x = self.transformer_encoder(x)
        # unlike the classifier model, every token is returned, so every row is normalised
        return F_.LayerNormFn.apply(x, self.layer_norm.weight, self.layer_norm.bias, self.layer_norm.eps)
