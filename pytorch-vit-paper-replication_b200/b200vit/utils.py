"""Checkpoint interoperability.

* State dicts written by the reference's `going_modular/utils.py:save_model` (`torch.save(model.state_dict())`,
  utils.py:7-35) load into the b200vit models as they are: the module tree and therefore the 152 keys are identical.
* The reference's best results come from torchvision's pretrained `vit_b_16` (08_..._self.ipynb:4039-4129), a model
  with the same mathematics but other attribute names and LayerNorm eps 1e-6. `load_torchvision_vit` copies such a
  state dict into a b200vit `ViT` (classifier or headless) and switches the LayerNorms to that eps.
"""
from __future__ import annotations

import re
from typing import Dict

import torch
from torch import nn

_LAYER = re.compile(r"^encoder\.layers\.encoder_layer_(\d+)\.(.+)$")
_IN_LAYER = {
    "ln_1.weight": "msa_block.layer_norm.weight", "ln_1.bias": "msa_block.layer_norm.bias",
    "self_attention.in_proj_weight": "msa_block.multi_head_attention.in_proj_weight",
    "self_attention.in_proj_bias": "msa_block.multi_head_attention.in_proj_bias",
    "self_attention.out_proj.weight": "msa_block.multi_head_attention.out_proj.weight",
    "self_attention.out_proj.bias": "msa_block.multi_head_attention.out_proj.bias",
    "ln_2.weight": "mlp_block.layer_norm.weight", "ln_2.bias": "mlp_block.layer_norm.bias",
    "mlp.0.weight": "mlp_block.mlp.0.weight", "mlp.0.bias": "mlp_block.mlp.0.bias",
    "mlp.3.weight": "mlp_block.mlp.3.weight", "mlp.3.bias": "mlp_block.mlp.3.bias",
    # older torchvision checkpoints
    "mlp.linear_1.weight": "mlp_block.mlp.0.weight", "mlp.linear_1.bias": "mlp_block.mlp.0.bias",
    "mlp.linear_2.weight": "mlp_block.mlp.3.weight", "mlp.linear_2.bias": "mlp_block.mlp.3.bias",
}
_TOP = {
    "class_token": "patch_embedding_block.class_token",
    "conv_proj.weight": "patch_embedding_block.patch_and_flatten.0.weight",
    "conv_proj.bias": "patch_embedding_block.patch_and_flatten.0.bias",
    "encoder.pos_embedding": "patch_embedding_block.position_embedding",
    "encoder.ln.weight": "layer_norm.weight", "encoder.ln.bias": "layer_norm.bias",
    "heads.head.weight": "classifier.0.weight", "heads.head.bias": "classifier.0.bias",
}


def torchvision_vit_key(key: str) -> str:
    """Name of a torchvision VisionTransformer state-dict entry in the reference's / b200vit's module tree."""
    if key in _TOP:
        return _TOP[key]
    m = _LAYER.match(key)
    if m and m.group(2) in _IN_LAYER:
        return f"transformer_encoder.{m.group(1)}.{_IN_LAYER[m.group(2)]}"
    raise KeyError(f"unexpected torchvision ViT key: {key}")


def load_torchvision_vit(model: nn.Module, tv_state_dict: Dict[str, torch.Tensor], layer_norm_eps: float = 1e-6,
                         load_head: bool = True) -> nn.Module:
    """Copy a torchvision `vit_b_16` / `vit_l_16` ... state dict into `model` (a b200vit ViT of matching geometry).

    `load_head=False` leaves the classifier alone (fine-tuning with a new head); a headless model ignores the head
    entries. All LayerNorms of `model` are set to `layer_norm_eps` (torchvision builds them with 1e-6; the reference
    and nn.LayerNorm default to 1e-5)."""
    own = model.state_dict()
    mapped = {}
    for k, v in tv_state_dict.items():
        name = torchvision_vit_key(k)
        if name.startswith("classifier.") and (not load_head or name not in own):
            continue
        if name not in own:
            raise KeyError(f"{k} -> {name}: no such parameter in the target model")
        if own[name].shape != v.shape:
            raise ValueError(f"{k} -> {name}: shape {tuple(v.shape)} does not match {tuple(own[name].shape)}")
        mapped[name] = v
    missing = [k for k in own if k not in mapped and not k.startswith("classifier.")]
    if missing:
        raise KeyError(f"torchvision state dict lacks {missing[:4]}{'...' if len(missing) > 4 else ''}")
    model.load_state_dict(mapped, strict=False)
    for mod in model.modules():
        if isinstance(mod, nn.LayerNorm):
            mod.eps = layer_norm_eps
    return model
