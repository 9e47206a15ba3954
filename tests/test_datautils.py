"""CPU: the data-side functions of the hot path are bit-exact against the reference fixtures."""
import torch

from building_llm_from_scratch_b200.datautils.dataloader_instruction_finetune import custom_collate_fn
from building_llm_from_scratch_b200.datautils.dataset import DatasetPT, shift_windows
from building_llm_from_scratch_b200.datautils.synthetic import alpaca_batch, gutenberg_batch


def test_collate_bit_exact(golden):
    for case in golden("data_pipeline.pt")["collate"]:
        x, y = custom_collate_fn(case["batch"], **case["kw"])
        assert x.dtype == torch.int64 and y.dtype == torch.int64
        assert torch.equal(x, case["inputs"]) and torch.equal(y, case["targets"]), case["batch"]


def test_collate_matches_oracle_on_random_batches():
    from oracle import reference_ops as R
    gen = torch.Generator().manual_seed(0)
    for _ in range(50):
        n = int(torch.randint(1, 8, (1,), generator=gen))
        batch = []
        for _ in range(n):
            ln = int(torch.randint(1, 60, (1,), generator=gen))
            toks = torch.randint(50250, 50262, (ln,), generator=gen).tolist()   # dense in pad ids
            batch.append((int(torch.randint(0, ln + 3, (1,), generator=gen)), toks))
        kw = {} if n % 2 else dict(allowed_max_length=int(torch.randint(1, 70, (1,), generator=gen)))
        a, b = custom_collate_fn(batch, **kw)
        c, d = R.collate(batch, **kw)
        assert torch.equal(a, c) and torch.equal(b, d)


def test_shift_bit_exact(golden):
    for case in golden("data_pipeline.pt")["shift"]:
        x, y = shift_windows(case["tokens"], case["max_length"], case["stride"])
        assert x.tolist() == case["inputs"] and y.tolist() == case["targets"]

    class Tok:
        def encode(self, txt, allowed_special=None):
            return list(range(100, 133))

    ds = DatasetPT("x", Tok(), 8, 8)
    assert len(ds) == 4 and ds[1][0].tolist() == list(range(108, 116)) and ds[1][1].tolist() == list(range(109, 117))


def test_synthetic_shapes():
    gen = torch.Generator().manual_seed(1)
    x, y = alpaca_batch(gen, 4, 256, 128256)
    assert x.shape == (4, 256) and y.shape == (4, 256) and x.dtype == torch.int64
    frac = float((y == -100).float().mean())
    assert 0.2 < frac < 0.95
    x, y = gutenberg_batch(gen, 3, 64, 50257)
    assert x.shape == (3, 64) and torch.equal(x[:, 1:], y[:, :-1]) and (y != -100).all()


def test_mixed_precision_policy_objects():
    from building_llm_from_scratch_b200.datautils import mixed_precision as mp
    assert set(mp.mixed_precision_policies) == {"fp16", "bf16", "bf16_hybrid", "fp32"}
    assert mp.bf16_hybrid_policy.param_dtype == torch.float32 and mp.bf16_hybrid_policy.reduce_dtype == torch.bfloat16
    assert mp.fpSixteen is mp.fp16_policy and mp.bfSixteen is mp.bf16_policy


def test_llama2_module_tree_and_state_dict_keys():
    from building_llm_from_scratch_b200.Models.Llama.Llama2 import Llama2Model
    cfg = dict(vocab_size=64, context_length=16, emb_dim=128, n_heads=2, n_layers=1, hidden_dim=64, dtype=torch.bfloat16)
    m = Llama2Model(cfg)
    keys = set(m.state_dict())
    assert {"tok_emb.weight", "trf_blocks.0.att.W_query.weight", "trf_blocks.0.att.mask", "trf_blocks.0.att.cos",
            "trf_blocks.0.att.sin", "trf_blocks.0.ff.fc3.weight", "trf_blocks.0.norm1.weight", "final_norm.weight",
            "out_head.weight"} <= keys
    assert m.state_dict()["trf_blocks.0.att.cos"].dtype == torch.float32
    assert m.load_state_dict(m.state_dict()).missing_keys == []
