"""b200vit -- B200-native (sm_100a) implementation of the ViT-B/16 training step of
AvalonEnjoyer/pytorch-ViT-paper-replication, behind the reference's own Python API.

    from b200vit.models.vit import ViT                   # reference: models/vit.py
    from b200vit.going_modular import engine             # reference: going_modular/going_modular/engine.py

All arithmetic runs in hand-written CUDA kernels reached through the C ABI of libb200vit.so
(include/b200vit.h). There is no CPU or PyTorch fallback.

    with b200vit.precision("fp32"): ...                  # strict mode: logits within 1e-3 of the fp32 reference
"""
__version__ = "0.2.0"

from .functional import get_precision, precision, set_precision  # noqa: E402,F401
