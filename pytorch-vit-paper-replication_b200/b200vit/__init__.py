"""b200vit -- B200-native (sm_100a) implementation of the ViT-B/16 training step of
AvalonEnjoyer/pytorch-ViT-paper-replication, behind the reference's own Python API.

    from b200vit.models.vit import ViT                   # reference: models/vit.py
    from b200vit.going_modular import engine             # reference: going_modular/going_modular/engine.py

All arithmetic runs in hand-written CUDA kernels reached through the C ABI of libb200vit.so
(include/b200vit.h). There is no CPU or PyTorch fallback.
"""
__version__ = "0.1.0"
