"""Drop-in for datautils/mixed_precision.py:10-46 — the four FSDP MixedPrecision policy objects and
the lookup dict, plus the two alias names build_components.py:166 tries to import (they do not
exist upstream). The FSDP consumer itself is out of scope; what the hot path needs
from a policy is the dtype casts, done here by the b200llm cast kernels."""
from __future__ import annotations

import torch
from torch.distributed.fsdp import MixedPrecision

fp16_policy = MixedPrecision(param_dtype=torch.float16, reduce_dtype=torch.float16, buffer_dtype=torch.float16)
bf16_policy = MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16)
bf16_hybrid_policy = MixedPrecision(param_dtype=torch.float32, reduce_dtype=torch.bfloat16,
                                    buffer_dtype=torch.bfloat16)
fp32_policy = MixedPrecision(param_dtype=torch.float32, reduce_dtype=torch.float32, buffer_dtype=torch.float32)

mixed_precision_policies = {"fp16": fp16_policy, "bf16": bf16_policy, "bf16_hybrid": bf16_hybrid_policy,
                            "fp32": fp32_policy}

fpSixteen = fp16_policy      # names imported by build_components.py:166
bfSixteen = bf16_policy


def cast_to_policy(t: torch.Tensor, policy: MixedPrecision, kind: str = "param") -> torch.Tensor:
    """Cast a CUDA tensor to the policy's param / reduce / buffer dtype with the b200llm cast kernels
    (fp32 <-> bf16); other dtype pairs are rejected (fp16 has no kernel on this path)."""
    from .. import kernels as K
    want = {"param": policy.param_dtype, "reduce": policy.reduce_dtype, "buffer": policy.buffer_dtype}[kind]
    if want is None or t.dtype == want:
        return t
    return K.cast(t, want)


def optimizer_kwargs(policy: MixedPrecision) -> dict:
    """What a policy means for optim.B200AdamW on this path (the FSDP consumer of these objects is broken upstream,
    SURVEY.md R6): parameters kept in fp32 while compute / reduction run in bf16 (bf16_hybrid, :24-28) = fp32 master
    weights and fp32 moments behind the bf16 compute copy; every other policy keeps optimizer state in the parameter
    dtype, as the reference's plain AdamW does."""
    hybrid = policy.param_dtype == torch.float32 and policy.reduce_dtype == torch.bfloat16
    return {"master_weights": bool(hybrid)}
