"""CPU: the cache of tensors derived from frozen weights (row concatenation for the packed projections,
128-row interleaving for the fused SwiGLU forward) follows the weight OBJECTS: a rewrite of a weight rebuilds
the entry, a dead weight removes it, and a new tensor that happens to reuse an address cannot hit a stale one."""
import gc

import torch

from building_llm_from_scratch_b200 import ops


def _frozen(*shape):
    return torch.nn.Parameter(torch.randn(*shape), requires_grad=False)


def test_concatenation_cache_tracks_versions_and_lifetimes():
    a, b = _frozen(4, 3), _frozen(2, 3)
    c1 = ops._packed_weight([a, b])
    assert torch.equal(c1, torch.cat([a, b])) and ops._packed_weight([a, b]) is c1
    with torch.no_grad():
        a.mul_(2.0)                                   # what load_state_dict does: in-place write, version bump
    c2 = ops._packed_weight([a, b])
    assert c2 is not c1 and torch.equal(c2, torch.cat([a, b]))
    n = len(ops._wcat_cache)
    del a, b
    gc.collect()
    assert len(ops._wcat_cache) == n - 1
    t = _frozen(4, 3)
    t.requires_grad_(True)
    assert ops._packed_weight([t, _frozen(2, 3)]) is None      # trainable weights are never cached


def test_interleaving_layout():
    w1, w2 = _frozen(256, 8), _frozen(256, 8)
    il = ops._interleaved_weight([w1, w2])
    assert il.shape == (512, 8)
    assert torch.equal(il[:128], w1[:128]) and torch.equal(il[128:256], w2[:128])
    assert torch.equal(il[256:384], w1[128:]) and torch.equal(il[384:], w2[128:])
    assert ops._interleaved_weight([_frozen(200, 8), _frozen(200, 8)]) is None     # not a multiple of 128 rows
    b1, b2 = torch.randn(8, 256), torch.randn(8, 256)
    bil = ops._interleaved_lora_b(b1, b2)
    assert bil.shape == (16, 512)
    assert torch.equal(bil[:8, :128], b1[:, :128]) and float(bil[:8, 128:256].abs().sum()) == 0.0
    assert torch.equal(bil[8:, 128:256], b2[:, :128]) and float(bil[8:, :128].abs().sum()) == 0.0
