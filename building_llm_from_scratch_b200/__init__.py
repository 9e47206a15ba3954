"""B200-native decoder training step for chemphenoms/Building_LLM_from_scratch (hot path only).

Layout: csrc/ (sm_100a kernels + C-ABI), _lib.py (ctypes binding), kernels.py (tensor wrappers),
ops.py (torch.library custom ops), Models/, lora.py, train.py (drop-in mirrors of the reference
interface for this path).
"""
__version__ = "0.1.0"
