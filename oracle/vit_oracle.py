"""CPU oracle for the ViT training-step hot path.  TEST INFRASTRUCTURE ONLY.

A plain fp32 restatement, in explicit tensor arithmetic, of what the reference computes on its hot path
(`models/vit.py` ViT.forward driven by `going_modular/going_modular/engine.py` train_step). The reference
delegates all arithmetic to torch.nn modules (third-party, torch unpinned, validated here against
torch 2.11.0); this file restates that arithmetic op by op so the CUDA path has an independent checker.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` leg may import this
module. The product package (b200vit) never does: it has no CPU fallback.

Parity pinning: `oracle/make_golden.py` imports the UNMODIFIED reference from /root/reference and writes
golden vectors (tests/golden/*.npz); tests/test_oracle.py checks this restatement against them, including the
reference notebook's known-answer logits [[0.6136, -0.9092, 0.5918]] (08_pytorch_paper_replicating_self.ipynb
:2605-2631). Parity is therefore pinned to the live reference at the torch version in this image.

State is a dict with the reference's state_dict keys (152 tensors for ViT-B/16).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

Tensor = torch.Tensor


@dataclass
class ViTConfig:
    """Constructor arguments of the reference ViT (models/vit.py:173-183), same names and defaults."""
    image_size: int = 224
    patch_size: int = 16
    num_transformer_layer: int = 12
    num_heads: int = 12
    embedding_dim: int = 768
    mlp_size: int = 3072
    attn_dropout: float = 0.0
    mlp_dropout: float = 0.1
    embedding_dropout: float = 0.1
    num_classes: Optional[int] = 1000  # None -> headless variant (models/vit_no_classifier.py)

    @property
    def tokens(self) -> int:
        return (self.image_size // self.patch_size) ** 2 + 1


# ---------------------------------------------------------------------------------------------------------
# forward pieces
# ---------------------------------------------------------------------------------------------------------
def layer_norm(x: Tensor, gamma: Tensor, beta: Tensor, eps: float = 1e-5) -> Tensor:
    """nn.LayerNorm(D): biased variance over the last dim (models/vit.py:83,115,216)."""
    mean = x.mean(dim=-1, keepdim=True)
    var = ((x - mean) ** 2).mean(dim=-1, keepdim=True)
    return (x - mean) / torch.sqrt(var + eps) * gamma + beta


def gelu(x: Tensor) -> Tensor:
    """nn.GELU() default = exact erf form (models/vit.py:121)."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def dropout(x: Tensor, p: float, training: bool, generator: Optional[torch.Generator] = None) -> Tensor:
    """nn.Dropout: zero with probability p, scale survivors by 1/(1-p) (models/vit.py:43,122,125)."""
    if not training or p <= 0.0:
        return x
    # same generator consumption as ATen's CPU native_dropout (a bernoulli_(1-p) mask), so that with the same
    # torch.manual_seed the oracle reproduces the reference's train-mode outputs, e.g. the notebook KAT
    keep = torch.empty_like(x).bernoulli_(1.0 - p, generator=generator)
    return x * keep / (1.0 - p)


def patch_embedding(img: Tensor, sd: Dict[str, Tensor], cfg: ViTConfig, training: bool,
                    prefix: str = "patch_embedding_block.") -> Tensor:
    """PatchEmbedding.forward (models/vit.py:45-67): stride==kernel conv as a patch GEMM, prepend the class
    token, add the position embedding, dropout."""
    b, c, h, w = img.shape
    p = cfg.patch_size
    wconv = sd[prefix + "patch_and_flatten.0.weight"]           # [D, C, P, P]
    bconv = sd[prefix + "patch_and_flatten.0.bias"]
    d = wconv.shape[0]
    # [B, C, H/P, P, W/P, P] -> [B, H/P, W/P, C, P, P] -> [B, Np, C*P*P]
    patches = img.reshape(b, c, h // p, p, w // p, p).permute(0, 2, 4, 1, 3, 5).reshape(b, (h // p) * (w // p), c * p * p)
    x = patches @ wconv.reshape(d, -1).t() + bconv               # [B, Np, D]
    cls = sd[prefix + "class_token"].expand(b, -1, -1)
    x = torch.cat((cls, x), dim=1) + sd[prefix + "position_embedding"]
    return dropout(x, cfg.embedding_dropout, training)


def self_attention(xn: Tensor, in_w: Tensor, in_b: Tensor, out_w: Tensor, out_b: Tensor, heads: int) -> Tensor:
    """nn.MultiheadAttention(batch_first=True) with q=k=v, need_weights=False, no mask, attention dropout 0
    (models/vit.py:86-97): packed in-projection rows are [q; k; v], heads are contiguous slices of D."""
    b, t, d = xn.shape
    hd = d // heads
    qkv = xn @ in_w.t() + in_b                                   # [B, T, 3D]
    q, k, v = qkv.split(d, dim=-1)
    q = q.reshape(b, t, heads, hd).transpose(1, 2)
    k = k.reshape(b, t, heads, hd).transpose(1, 2)
    v = v.reshape(b, t, heads, hd).transpose(1, 2)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
    pattn = torch.softmax(s, dim=-1)
    o = (pattn @ v).transpose(1, 2).reshape(b, t, d)
    return o @ out_w.t() + out_b


def msa_block(x: Tensor, sd: Dict[str, Tensor], prefix: str, heads: int) -> Tensor:
    """MultiHeadSelfAttentionBlock.forward (models/vit.py:92-98): LN then self-attention, no residual."""
    xn = layer_norm(x, sd[prefix + "layer_norm.weight"], sd[prefix + "layer_norm.bias"])
    a = prefix + "multi_head_attention."
    return self_attention(xn, sd[a + "in_proj_weight"], sd[a + "in_proj_bias"], sd[a + "out_proj.weight"],
                          sd[a + "out_proj.bias"], heads)


def mlp_block(x: Tensor, sd: Dict[str, Tensor], prefix: str, p: float, training: bool) -> Tensor:
    """MLPBlock.forward (models/vit.py:128-131): LN -> Linear -> GELU -> Dropout -> Linear -> Dropout."""
    xn = layer_norm(x, sd[prefix + "layer_norm.weight"], sd[prefix + "layer_norm.bias"])
    h = gelu(xn @ sd[prefix + "mlp.0.weight"].t() + sd[prefix + "mlp.0.bias"])
    h = dropout(h, p, training)
    y = h @ sd[prefix + "mlp.3.weight"].t() + sd[prefix + "mlp.3.bias"]
    return dropout(y, p, training)


def encoder_block(x: Tensor, sd: Dict[str, Tensor], prefix: str, cfg: ViTConfig, training: bool) -> Tensor:
    """TransformerEncoderBlock.forward (models/vit.py:166-169): pre-norm residuals."""
    x = msa_block(x, sd, prefix + "msa_block.", cfg.num_heads) + x
    x = mlp_block(x, sd, prefix + "mlp_block.", cfg.mlp_dropout, training) + x
    return x


def vit_forward(img: Tensor, sd: Dict[str, Tensor], cfg: ViTConfig, training: bool = False) -> Tensor:
    """ViT.forward (models/vit.py:224-236); headless variant returns the normalised sequence
    (models/vit_no_classifier.py:217-226)."""
    x = patch_embedding(img, sd, cfg, training)
    for i in range(cfg.num_transformer_layer):
        x = encoder_block(x, sd, f"transformer_encoder.{i}.", cfg, training)
    x = layer_norm(x, sd["layer_norm.weight"], sd["layer_norm.bias"])
    if cfg.num_classes is None:
        return x
    return x[:, 0] @ sd["classifier.0.weight"].t() + sd["classifier.0.bias"]


# ---------------------------------------------------------------------------------------------------------
# loss / metrics / optimizer pieces of engine.train_step
# ---------------------------------------------------------------------------------------------------------
def cross_entropy(logits: Tensor, y: Tensor) -> Tensor:
    """nn.CrossEntropyLoss() with mean reduction (engine.py:53)."""
    z = logits - logits.max(dim=1, keepdim=True).values
    lse = torch.log(torch.exp(z).sum(dim=1))
    return (lse - z[torch.arange(len(y)), y]).mean()


def accuracy(logits: Tensor, y: Tensor) -> float:
    """argmax(softmax(logits)) == y, mean over the batch (engine.py:73-74)."""
    return (logits.argmax(dim=1) == y).sum().item() / len(y)


def clip_coefficient(grads: Sequence[Tensor], max_norm: float = 1.0) -> Tuple[float, float]:
    """torch.nn.utils.clip_grad_norm_ (engine.py:63): returns (total_norm, coef = min(1, max_norm/(norm+1e-6)))."""
    total = math.sqrt(sum(float((g.double() ** 2).sum()) for g in grads))
    return total, min(1.0, max_norm / (total + 1e-6))


def adam_update(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, lr: float, weight_decay: float,
                beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, decoupled: bool = False) -> None:
    """torch.optim.Adam single-tensor update, in place. decoupled=False is the reference recipe
    (`decoupled_weight_decay: False`, notebook :2862): L2 decay added to the gradient."""
    if decoupled:
        p.mul_(1.0 - lr * weight_decay)
    elif weight_decay != 0.0:
        g = g + weight_decay * p
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    p.addcdiv_(m, denom, value=-lr / bc1)


def lr_at(step: int, total_steps: int, base_lr: float = 1e-3, warmup_frac: float = 0.05,
          start_factor: float = 1e-6) -> float:
    """Learning rate used by optimizer step number `step` (0-based) under the notebook schedule
    (:2941-2960): SequentialLR([LinearLR(start_factor -> 1 over `warmup` steps), LinearLR(1 -> 0 over the rest)],
    milestones=[warmup]) stepped once per batch after optimizer.step() (engine.py:66-68)."""
    warmup = int(warmup_frac * total_steps)
    if step < warmup:
        return base_lr * (start_factor + (1.0 - start_factor) * step / warmup)
    decay_steps = total_steps - warmup
    k = min(step - warmup, decay_steps)
    return base_lr * (1.0 + (0.0 - 1.0) * k / decay_steps)


def decay_split(names_and_params) -> Tuple[List[str], List[str]]:
    """Weight-decay split of the notebook (:2791-2802): ndim == 1 or name endswith '.bias' -> no decay."""
    decay, no_decay = [], []
    for name, p in names_and_params:
        (no_decay if (p.ndim == 1 or name.endswith(".bias")) else decay).append(name)
    return decay, no_decay


class OracleTrainer:
    """engine.train_step restated for one model held as a state dict (engine.py:9-79) with the notebook recipe:
    CrossEntropyLoss, clip_grad_norm_(1.0), Adam(lr, coupled weight decay 0.03 on matrices / 0 elsewhere),
    per-batch warm-up + linear-decay schedule. Autograd supplies the backward pass; everything else is the
    explicit arithmetic above. Used as the numerical oracle for loss curves and as the timed CPU baseline."""

    def __init__(self, sd: Dict[str, Tensor], cfg: ViTConfig, total_steps: int, base_lr: float = 1e-3,
                 weight_decay: float = 0.03, max_norm: float = 1.0, scheduled: bool = True):
        self.cfg = cfg
        self.sd = {k: v.detach().clone().float().requires_grad_(True) for k, v in sd.items()}
        self.m = {k: torch.zeros_like(v) for k, v in self.sd.items()}
        self.v = {k: torch.zeros_like(v) for k, v in self.sd.items()}
        decay, _ = decay_split(self.sd.items())
        self.decay = set(decay)
        self.total_steps = total_steps
        self.base_lr = base_lr
        self.weight_decay = weight_decay
        self.max_norm = max_norm
        self.scheduled = scheduled
        self.step_idx = 0

    def step(self, X: Tensor, y: Tensor, training_dropout: bool = False) -> Tuple[float, float]:
        logits = vit_forward(X, self.sd, self.cfg, training=training_dropout)
        loss = cross_entropy(logits, y)
        names = list(self.sd.keys())
        grads = torch.autograd.grad(loss, [self.sd[k] for k in names])
        _, coef = clip_coefficient(grads, self.max_norm)
        lr = lr_at(self.step_idx, self.total_steps, self.base_lr) if self.scheduled else self.base_lr
        with torch.no_grad():
            for k, g in zip(names, grads):
                adam_update(self.sd[k], g * coef, self.m[k], self.v[k], self.step_idx + 1, lr,
                            self.weight_decay if k in self.decay else 0.0)
        self.step_idx += 1
        return loss.item(), accuracy(logits.detach(), y)

    def epoch(self, loader) -> Tuple[float, float]:
        """engine.train_step over a whole loader (engine.py:45-79): mean of the per-batch losses / accuracies."""
        tot_l = tot_a = 0.0
        for X, y in loader:
            l, a = self.step(X, y)
            tot_l += l
            tot_a += a
        return tot_l / len(loader), tot_a / len(loader)


def evaluate(sd: Dict[str, Tensor], cfg: ViTConfig, loader) -> Tuple[float, float]:
    """engine.test_step (engine.py:81-130): eval-mode forward under no_grad, mean of per-batch loss / accuracy."""
    tot_l = tot_a = 0.0
    with torch.no_grad():
        for X, y in loader:
            logits = vit_forward(X, sd, cfg, training=False)
            tot_l += cross_entropy(logits, y).item()
            tot_a += accuracy(logits, y)
    return tot_l / len(loader), tot_a / len(loader)


def train(sd: Dict[str, Tensor], cfg: ViTConfig, train_loader, test_loader, epochs: int, total_steps: int):
    """engine.train (engine.py:132-211) with the notebook recipe: the four per-epoch metric lists."""
    tr = OracleTrainer(sd, cfg, total_steps=total_steps)
    res = {"train_loss": [], "train_acc": [], "test_loss": [], "test_acc": []}
    for _ in range(epochs):
        tl, ta = tr.epoch(train_loader)
        el, ea = evaluate(tr.sd, cfg, test_loader)
        res["train_loss"].append(tl); res["train_acc"].append(ta)
        res["test_loss"].append(el); res["test_acc"].append(ea)
    return res
