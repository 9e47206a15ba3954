"""GPU: the on-device collate / window kernels are bit-identical to the host drop-ins, which the CPU
suite pins against the reference's own outputs (tests/golden/collate.pt)."""
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(120)]


def test_device_collate_bit_exact_on_known_answers_and_random_batches(golden):
    from building_llm_from_scratch_b200.datautils.dataloader_instruction_finetune import custom_collate_fn
    from building_llm_from_scratch_b200.datautils.device_collate import custom_collate_fn_device
    P = 50256
    cases = [
        ([(2, [10, 11, 12, 13, 14]), (3, [20, 21, 22]), (1, [30])], None),
        ([(2, [10, P, 12, P, 14]), (1, [20, 21])], None),                       # natural pad ids inside the text
        ([(2, list(range(10, 17))), (3, [20, 21, 22])], 4),                     # allowed_max_length truncation
        ([(0, [10, 11, 12]), (1, [20, 21, 22])], None),                         # instruction_length 0 and 1
        ([(1, [10, 11, 12])], None),
    ]
    g = torch.Generator().manual_seed(0)
    for _ in range(20):
        n = int(torch.randint(1, 9, (1,), generator=g))
        batch = []
        for _ in range(n):
            ln = int(torch.randint(1, 700, (1,), generator=g))
            toks = torch.randint(0, 128256, (ln,), generator=g)
            toks[torch.rand(ln, generator=g) < 0.02] = P                        # sprinkle natural pad ids
            batch.append((int(torch.randint(0, ln + 3, (1,), generator=g)), toks.tolist()))
        cases.append((batch, [None, 64, 300, 2048][int(torch.randint(0, 4, (1,), generator=g))]))
    for batch, amax in cases:
        hi, ht = custom_collate_fn(batch, allowed_max_length=amax)
        di, dt = custom_collate_fn_device(batch, allowed_max_length=amax)
        assert di.dtype == torch.int64 and dt.dtype == torch.int64
        assert torch.equal(di.cpu(), hi) and torch.equal(dt.cpu(), ht), (batch[0][0], amax)
    # the first known answer, literally
    di, dt = custom_collate_fn_device(cases[0][0])
    assert dt.cpu().tolist() == [[-100, 12, 13, 14, P], [-100, -100, P, -100, -100], [P, -100, -100, -100, -100]]


def test_device_windows_match_dataset_shift():
    from building_llm_from_scratch_b200.datautils.dataset import shift_windows
    from building_llm_from_scratch_b200.datautils.device_collate import window_batch
    tok = torch.randint(0, 50257, (5000,), generator=torch.Generator().manual_seed(1))
    for L, stride in [(4, 4), (128, 128), (256, 100)]:
        hi, ht = shift_windows(tok, L, stride)
        starts = list(range(0, tok.numel() - L, stride))
        di, dt = window_batch(tok.cuda(), starts, L)
        assert torch.equal(di.cpu(), hi) and torch.equal(dt.cpu(), ht)
    with pytest.raises(ValueError):
        window_batch(tok.cuda(), [tok.numel() - 4], 4)
