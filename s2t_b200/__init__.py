"""s2t_b200 — B200-native (sm_100a) implementation of the xuchennlp/S2T speech-encoder training hot path.

Python/PyTorch host code -> ctypes -> ``libs2t_b200.so`` (hand-written CUDA: tcgen05/TMEM/TMA GEMMs,
fbank, LayerNorm, softmax, CTC).  No CPU fallback: importing is cheap, calling without the built library
or without a CUDA device raises.
"""
__version__ = "0.1.0"
