"""peft_b200 — B200-native (sm_100a) hot path for parameter-efficient fine-tuning of Vision Transformers.

Drop-in replacement for the nn.Module forward/backward of the reference's PEFT ViT blocks
(jkooy/Parameter-efficient-Fine-tuning-for-Vision-Transformers): same module classes, parameter names and trainer
contract, arithmetic in hand-written CUDA behind the C ABI of include/peft_b200.h (csrc/libpeftb200.so).
There is no CPU fallback: using any op without the built library / a CUDA device raises.
"""
__version__ = "0.1.0"
