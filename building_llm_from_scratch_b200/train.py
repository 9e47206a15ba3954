"""Drop-in for the hot-path part of train.py: Trainer with the reference's constructor,
attributes, LR schedule (train.py:100-107) and step order (zero_grad, lr, forward + CE, backward,
clip 1.0, optimizer step; train.py:94-123). Evaluation / sampling / checkpoint orchestration is
kept behaviour-compatible (calc_loss_loader / evaluate_model / save_checkpoint below, generate.py) but
not separately accelerated beyond running on the same kernels; train_epoch /
train_model / finetune_model / generate_and_print_sample follow train.py:128-229 so the reference's main.py can
drive this class unchanged.
"""
from __future__ import annotations

import math

import torch

from . import ops
from .ddp import B200DDP
from .optim import B200AdamW


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class Trainer:
    def __init__(self, model, optimizer, config, data_files, loaderObj, save_dir, warmup_steps=10, initial_lr=1e-5,
                 min_lr=1e-6, device="cpu", rank=0, eval_freq=1, save_ckpt_freq=1, print_sample_iter=1, eval_iter=1):
        self.config = config
        self.model = model
        self.optimizer = optimizer
        self.data_files = data_files
        self.loaderObj = loaderObj
        self.save_dir = save_dir
        self.device = device
        self.rank = rank
        self.warmup_steps = warmup_steps
        self.initial_lr = initial_lr
        self.min_lr = min_lr
        self.eval_freq = eval_freq
        self.save_ckpt_freq = save_ckpt_freq
        self.print_sample_iter = print_sample_iter
        self.eval_iter = eval_iter
        self.global_step = -1
        self.tokens_seen = 0
        self.train_losses = []
        self.val_losses = []
        self.track_lrs = []
        self.track_tokens_seen = []
        self.last_loss = None
        if isinstance(optimizer, B200AdamW):
            optimizer.max_norm = 1.0  # train.py:115-120

    # train.py:154-156
    def set_schedule(self, total_training_steps, peak_lr=None):
        self.peak_lr = self.optimizer.param_groups[0]["lr"] if peak_lr is None else peak_lr
        self.total_training_steps = total_training_steps
        self.lr_increment = (self.peak_lr - self.initial_lr) / self.warmup_steps

    def lr_for_step(self, step):
        if step < self.warmup_steps:
            return self.initial_lr + step * self.lr_increment
        progress = (step - self.warmup_steps) / (self.total_training_steps - self.warmup_steps)
        return self.min_lr + (self.peak_lr - self.min_lr) * 0.5 * (1 + math.cos(math.pi * progress))

    # train.py:88-92
    def calc_loss_batch(self, input_batch, target_batch):
        target_batch_host = target_batch if not target_batch.is_cuda else None
        if target_batch_host is not None and isinstance(self.config, dict) and "vocab_size" in self.config:
            # F.cross_entropy refuses class indices outside [0, V) (other than ignore_index); checked here on the
            # host copy of the batch — targets that only exist on the device turn the loss into NaN instead
            V = self.config["vocab_size"]
            bad = (target_batch_host != -100) & ((target_batch_host < 0) | (target_batch_host >= V))
            if bool(bad.any()):
                raise IndexError(f"Target {int(target_batch_host[bad][0])} is out of bounds.")
        input_batch = input_batch.to(self.device, non_blocking=True)
        target_batch = target_batch.to(self.device, non_blocking=True)
        core = self.model.module if isinstance(self.model, B200DDP) else self.model
        if hasattr(core, "forward_loss"):   # fused vocab projection + CE, in training and in evaluation
            loss = core.forward_loss(input_batch, target_batch, capacity=self._head_capacity(target_batch_host))
        else:
            logits = self.model(input_batch)
            loss = ops.cross_entropy(logits, target_batch)[0]
        return loss.to(torch.bfloat16) if self._loss_bf16(core) else loss

    @staticmethod
    def _head_capacity(targets_host, granule=128):
        """Rows the vocab projection has to see: the non-ignored targets (counted on the host copy of
        the batch, no device sync), rounded up; 0 = all rows (targets already on the device)."""
        if targets_host is None:
            return 0
        n = int((targets_host != -100).sum())
        if n == 0 or n >= targets_host.numel():
            return 0
        return (n + granule - 1) // granule * granule

    @staticmethod
    def _loss_bf16(core):
        p = next(core.parameters())
        return p.dtype == torch.bfloat16

    # train.py:94-126 (exceptions propagate instead of being logged and swallowed)
    def train_batch(self, input_batch, target_batch):
        if not hasattr(self, "peak_lr"):
            raise RuntimeError("Trainer.train_batch needs the LR schedule: call set_schedule(total_training_steps) — "
                               "train_model / finetune_model do it, as train.py:154-156 does")
        self.optimizer.zero_grad()
        self.global_step += 1
        lr = self.lr_for_step(self.global_step)
        for param_group in self.optimizer.param_groups:
            param_group["lr"] = lr
        self.track_lrs.append(lr)
        loss = self.calc_loss_batch(input_batch, target_batch)
        loss.backward()
        fused = isinstance(self.optimizer, B200AdamW)
        if isinstance(self.model, B200DDP):
            if fused and self.optimizer.zero:
                pass          # ZeRO-1: optimizer.step() reduce-scatters the flat gradients (optim.py)
            elif fused:
                self.optimizer._absorb_autograd_grads()
                self.model.reduce_gradients(self.optimizer.flat_grads)
            else:
                self.model.reduce_gradients([p.grad for p in self.model.parameters() if p.grad is not None])
        if not fused:
            core = self.model.module if isinstance(self.model, B200DDP) else self.model
            torch.nn.utils.clip_grad_norm_(core.parameters(), max_norm=1.0)
        self.optimizer.step()
        self.tokens_seen += input_batch.numel()
        self.last_loss = loss.detach()
        return self.last_loss

    # ---- orchestration around the hot loop (train.py:128-229): same cadence and return values as the reference;
    # thin Python over train_batch / evaluate_model / generate, none of it on the timed path -------------------
    ANTONYM_PROMPT = ("Below is an instruction that describes a task. Write a response that appropriately completes the "
                      "request.\n\n### Instruction:\nWhat is an antonym of 'complicated'?")

    def train_epoch(self, epoch_no, train_loader, val_loader, start_context="Every effort moves you"):
        """train.py:128-151: one pass over train_loader; every eval_freq / print_sample_iter / save_ckpt_freq steps it
        evaluates, samples and checkpoints. Errors propagate (the reference logs and swallows them)."""
        self.model.train()
        for input_batch, target_batch in train_loader:
            self.train_batch(input_batch, target_batch)
            if self.global_step % self.eval_freq == 0:
                train_loss, val_loss = self.evaluate_model(train_loader, val_loader, self.eval_iter)
                self.train_losses.append(train_loss)
                self.val_losses.append(val_loss)
                self.track_tokens_seen.append(self.tokens_seen)
                if self.rank == 0:
                    print(f"Epoch {epoch_no + 1} | Step {self.global_step} | Train Loss: {train_loss:.3f} | "
                          f"Val Loss: {val_loss:.3f}")
            if self.global_step % self.print_sample_iter == 0:
                self.generate_and_print_sample(start_context)
            if self.global_step % self.save_ckpt_freq == 0:
                self.save_checkpoint(f"model_pg_{self.global_step}.pth")

    def _run_epochs(self, n_epochs, read, start_context):
        # train.py:154-156 — the schedule spans every file of every epoch
        self.set_schedule(self.loaderObj.get_total_steps_epoch(self.data_files) * n_epochs)
        try:
            for epoch in range(n_epochs):
                for file_path in self.data_files:
                    train_loader, val_loader = self.loaderObj.create_dataloaders(read(file_path), num_workers=2)
                    if hasattr(train_loader.sampler, "set_epoch"):
                        train_loader.sampler.set_epoch(epoch)
                    if start_context is None:
                        self.train_epoch(epoch, train_loader, val_loader)
                    else:
                        self.train_epoch(epoch, train_loader, val_loader, start_context=start_context)
        except KeyboardInterrupt:
            self.save_checkpoint(f"model_pg_{self.global_step}_interrupted.pth")
        return self.train_losses, self.val_losses, self.track_tokens_seen, self.track_lrs

    def train_model(self, n_epochs):
        """train.py:153-180: pre-training over text files (each gets the end-of-text marker appended)."""
        def read(path):
            with open(path, "r", encoding="utf-8") as f:
                return f.read() + " " + self.config["eos_text"] + " "
        return self._run_epochs(n_epochs, read, None)

    def finetune_model(self, n_epochs):
        """train.py:182-211: instruction fine-tuning over JSON files."""
        def read(path):
            import json
            with open(path, "r", encoding="utf-8") as f:
                return json.load(f)
        return self._run_epochs(n_epochs, read, self.ANTONYM_PROMPT)

    def generate_and_print_sample(self, start_context, temperature=1.0, top_k=5, memory_check=True, max_new_tokens=200):
        """train.py:213-229: sample max_new_tokens from the model (KV-cached here) and log the text on rank 0."""
        from .generate import generate
        core = self.model.module if isinstance(self.model, B200DDP) else self.model
        tok = self.loaderObj.tokenizer
        encoded = torch.tensor(tok.encode(start_context, allowed_special={self.config["eos_text"]})).unsqueeze(0)
        core.eval()
        token_ids = generate(model=core, idx=encoded.to(self.device), max_new_tokens=max_new_tokens,
                             context_size=self.config["context_length"], temperature=temperature, top_k=top_k,
                             eos_id=self.config["eos_id"])
        core.train()
        decoded = tok.decode(token_ids.squeeze(0).tolist())
        if self.rank == 0 and memory_check:
            print("Generated Sample: " + decoded.replace(chr(10), " "))
        return decoded

    # train.py:259-268
    def calc_loss_loader(self, data_loader, num_batches=None):
        total_loss = 0.0
        if len(data_loader) == 0:
            return float("nan")
        num_batches = min(num_batches or len(data_loader), len(data_loader))
        for i, (input_batch, target_batch) in enumerate(data_loader):
            if i >= num_batches:
                break              # the reference keeps iterating the whole loader; no need to
            total_loss += float(self.calc_loss_batch(input_batch, target_batch))
        return total_loss / num_batches

    # train.py:270-276
    def evaluate_model(self, train_loader, val_loader, eval_iter=5):
        self.model.eval()
        with torch.no_grad():
            train_loss = self.calc_loss_loader(train_loader, eval_iter)
            val_loss = self.calc_loss_loader(val_loader, eval_iter)
        self.model.train()
        return train_loss, val_loss

    # train.py:231-257 (state_dict of the unwrapped model, rank 0 only)
    def save_checkpoint(self, file_name):
        import os
        core = self.model.module if isinstance(self.model, B200DDP) else self.model
        if torch.distributed.is_initialized():
            torch.distributed.barrier()
        if self.rank == 0:
            torch.save(core.state_dict(), os.path.join(str(self.save_dir), file_name))
        if torch.distributed.is_initialized():
            torch.distributed.barrier()

    # Not in the reference (its checkpoints hold the model only and are never read back):
    # everything a resume needs, written off the training stream.
    def save_training_state(self, file_name, asynchronous=True):
        """Model state_dict (reference format) + optimizer moments / step + trainer counters. The device ->
        pinned-host copies run on a side stream and the file is written by a background thread, so the
        training stream only waits for the snapshot copies to be ENQUEUED; call .join() on the returned
        thread (or save with asynchronous=False) before relying on the file."""
        import os
        import threading
        core = self.model.module if isinstance(self.model, B200DDP) else self.model
        side = torch.cuda.Stream() if torch.cuda.is_available() else None
        snap = {}

        def to_host(t):
            if not t.is_cuda:
                return t.detach().clone()
            h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
            h.copy_(t.detach(), non_blocking=True)
            return h
        if side is not None:
            side.wait_stream(torch.cuda.current_stream())
        ctx = torch.cuda.stream(side) if side is not None else _null()
        with ctx:
            snap["model"] = {k: to_host(v) for k, v in core.state_dict().items()}
            snap["optimizer"] = self.optimizer.state_dict()     # B200AdamW: host copies of the flat moments
        snap["trainer"] = dict(global_step=self.global_step, tokens_seen=self.tokens_seen, track_lrs=list(self.track_lrs),
                               peak_lr=getattr(self, "peak_lr", None),
                               total_training_steps=getattr(self, "total_training_steps", None))
        done = torch.cuda.Event() if side is not None else None
        if done is not None:
            done.record(side)
            torch.cuda.current_stream().wait_event(done)   # later steps must not overwrite what is being copied
        path = os.path.join(str(self.save_dir), file_name)

        def write():
            if done is not None:
                done.synchronize()
            if self.rank == 0:
                torch.save(snap, path)
        th = threading.Thread(target=write, daemon=False)
        th.start()
        if not asynchronous:
            th.join()
        return th

    def load_training_state(self, file_name):
        import os
        snap = torch.load(os.path.join(str(self.save_dir), file_name), weights_only=False)
        core = self.model.module if isinstance(self.model, B200DDP) else self.model
        core.load_state_dict(snap["model"])
        self.optimizer.load_state_dict(snap["optimizer"])
        tr = snap["trainer"]
        self.global_step, self.tokens_seen, self.track_lrs = tr["global_step"], tr["tokens_seen"], list(tr["track_lrs"])
        if tr.get("total_training_steps") is not None:
            self.set_schedule(tr["total_training_steps"], tr["peak_lr"])
