"""CPU: the orchestration methods of the Trainer drop-in (train.py:128-229) keep the reference's cadence —
schedule from get_total_steps_epoch x epochs, evaluation / sampling / checkpoint every eval_freq /
print_sample_iter / save_ckpt_freq steps, the four tracking lists returned — driven here with stub loaders, a stub
tokenizer and a small CPU model that offers the same forward_loss / forward interface as the b200 models."""
import json

import torch


class _Model(torch.nn.Module):
    def __init__(self, V=32, d=8):
        super().__init__()
        self.emb = torch.nn.Embedding(V, d)
        self.head = torch.nn.Linear(d, V)

    def forward(self, x):
        return self.head(self.emb(x))

    def forward_loss(self, x, y, capacity=0):
        return torch.nn.functional.cross_entropy(self(x).flatten(0, 1), y.flatten())


class _Loader(list):
    class _S:
        def __init__(self):
            self.epochs = []

        def set_epoch(self, e):
            self.epochs.append(e)

    def __init__(self, items):
        super().__init__(items)
        self.sampler = self._S()


class _Tok:
    def encode(self, text, allowed_special=None):
        return [ord(c) % 32 for c in text][:6]

    def decode(self, ids):
        return " ".join(str(i) for i in ids)


class _LoaderObj:
    tokenizer = _Tok()

    def __init__(self):
        self.created, self.loaders = [], []

    def get_total_steps_epoch(self, files):
        return 3 * len(files)

    def create_dataloaders(self, data, num_workers=0):
        self.created.append(data)
        g = torch.Generator().manual_seed(len(self.created))
        mk = lambda n: _Loader([(torch.randint(0, 32, (2, 5), generator=g), torch.randint(0, 32, (2, 5), generator=g)) for _ in range(n)])
        tl, vl = mk(3), mk(2)
        self.loaders.append(tl)
        return tl, vl


def _trainer(tmp_path, files, **kw):
    from building_llm_from_scratch_b200.train import Trainer
    torch.manual_seed(0)
    model = _Model()
    opt = torch.optim.AdamW(model.parameters(), lr=5e-3, weight_decay=0.1)
    cfg = dict(vocab_size=32, context_length=16, eos_text="<|endoftext|>", eos_id=31)
    lo = _LoaderObj()
    return Trainer(model, opt, cfg, files, lo, tmp_path, warmup_steps=2, **kw), lo, model


def test_train_model_cadence_and_schedule(tmp_path):
    f1, f2 = tmp_path / "a.txt", tmp_path / "b.txt"
    f1.write_text("alpha")
    f2.write_text("beta")
    tr, lo, model = _trainer(tmp_path, [f1, f2], eval_freq=2, save_ckpt_freq=4, print_sample_iter=100, eval_iter=1)
    out = tr.train_model(n_epochs=2)
    assert tr.total_training_steps == 3 * 2 * 2 and tr.peak_lr == 5e-3            # train.py:154-156
    assert tr.global_step == 11 and tr.tokens_seen == 12 * 10
    assert lo.created == ["alpha <|endoftext|> ", "beta <|endoftext|> "] * 2     # text + end-of-text marker per file
    assert [l.sampler.epochs for l in lo.loaders] == [[0], [0], [1], [1]]
    assert len(tr.train_losses) == len(tr.val_losses) == len(tr.track_tokens_seen) == 6      # steps 0,2,...,10
    assert len(tr.track_lrs) == 12 and tr.track_lrs[0] == 1e-5 and abs(tr.track_lrs[2] - 5e-3) < 1e-12
    assert out == (tr.train_losses, tr.val_losses, tr.track_tokens_seen, tr.track_lrs)
    assert sorted(p.name for p in tmp_path.glob("model_pg_*.pth")) == ["model_pg_0.pth", "model_pg_4.pth", "model_pg_8.pth"]
    sd = torch.load(tmp_path / "model_pg_8.pth")
    assert set(sd) == set(model.state_dict())


def test_finetune_model_reads_json_and_samples(tmp_path, capsys):
    f = tmp_path / "d.json"
    f.write_text(json.dumps([{"instruction": "x", "output": "y"}]))
    tr, lo, _ = _trainer(tmp_path, [f], eval_freq=100, save_ckpt_freq=100, print_sample_iter=2, eval_iter=1)
    tr.finetune_model(n_epochs=1)
    assert lo.created == [[{"instruction": "x", "output": "y"}]]
    assert capsys.readouterr().out.count("Generated Sample:") == 2                 # steps 0 and 2
    assert tr.model.training


def test_train_batch_needs_a_schedule(tmp_path):
    import pytest
    tr, _, _ = _trainer(tmp_path, [])
    with pytest.raises(RuntimeError, match="set_schedule"):
        tr.train_batch(torch.zeros(1, 4, dtype=torch.long), torch.zeros(1, 4, dtype=torch.long))
