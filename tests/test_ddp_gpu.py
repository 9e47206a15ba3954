"""GPU, world_size 2 over NCCL (skipped on a single-GPU box): after a few TrainStep steps on different
per-rank batches every rank must hold bit-identical parameters — the all-reduce result is identical on
all ranks, so any gradient that escaped the exchange (e.g. a weight-gradient GEMM launched after it)
shows up as divergence. Covers the two-graph replay, captured collectives with overlapped buckets, and
eager steps."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, mode, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["B200_GRAPH_NCCL"] = "1" if mode == "graph_nccl" else "0"
    zero = mode.startswith("zero")
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.ddp import B200DDP
    from building_llm_from_scratch_b200.train import Trainer
    from building_llm_from_scratch_b200.engine import TrainStep
    cfg = dict(vocab_size=1024, context_length=256, emb_dim=256, n_heads=4, n_layers=3, hidden_dim=512,
               n_kv_groups=2, rope_base=10000.0, rope_freq=None, dtype=torch.bfloat16)
    torch.manual_seed(123 + rank)                      # ranks start different: the wrapper broadcasts rank 0's
    dev = torch.device("cuda", rank)
    model = Llama3Model(cfg).to(dev)
    for p in model.parameters():
        p.requires_grad_(False)
    replace_linear_with_lora(model, rank=16, alpha=32, dtype=torch.bfloat16)
    model.to(dev)
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.endswith("lora.B"):
                p.normal_(0, 0.02)
    ddp = B200DDP(model, bucket_mb=0.05)
    opt = B200AdamW(model.parameters(), lr=1e-3, weight_decay=0.1, zero=zero)
    tr = Trainer(ddp, opt, cfg, [], None, "/tmp", device=dev, warmup_steps=2)
    tr.set_schedule(total_training_steps=100, peak_lr=1e-3)
    step = TrainStep(tr, (2, 256), use_graph=(mode not in ("eager", "zero_eager")))
    step.compact_head = False
    step.capture(warmup=2)
    g = torch.Generator().manual_seed(1000 + rank)
    for _ in range(4):
        x = torch.randint(0, 1024, (2, 256), generator=g)
        y = torch.randint(0, 1024, (2, 256), generator=g)
        y[rank, :100 * (rank + 1)] = -100
        step.run(x.to(dev), y.to(dev))
    torch.cuda.synchronize()
    flat = torch.cat([g_["p"].float().reshape(-1) for g_ in opt._groups.values()])
    gathered = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    if rank == 0:
        n_state = sum(g_["m"].numel() for g_ in opt._groups.values())
        out.put(dict(equal=bool(torch.equal(gathered[0], gathered[1])), params=gathered[0].cpu().numpy(), n_state=n_state,
                     n_params=flat.numel(),
                     maxdiff=float((gathered[0] - gathered[1]).abs().max()),
                     moved=float((flat - flat.mean()).abs().sum()), loss=float(step.loss)))
    step.close()
    dist.barrier()
    dist.destroy_process_group()


def _run(mode):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, out)) for r in range(2)]
    for p in procs:
        p.start()
    # read the result BEFORE joining: a payload larger than the pipe buffer blocks the writer until it is read
    import time
    res, t0 = None, time.time()
    while res is None and time.time() - t0 < 150:
        if not out.empty():
            res = out.get()
        elif any(p.exitcode not in (None, 0) for p in procs):
            break
        else:
            time.sleep(0.2)
    for p in procs:
        p.join(timeout=60)
        if p.exitcode is None:
            p.kill()
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    assert res is not None, "no result from rank 0"
    return res


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("mode", ["two_graphs", "graph_nccl", "eager"])
def test_ranks_stay_bit_identical(mode):
    res = _run(mode)
    assert res["equal"], f"parameters diverged across ranks by {res['maxdiff']} ({mode})"
    assert res["loss"] == res["loss"]      # not NaN


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_zero1_matches_replicated_optimizer():
    """ZeRO-1 (optimizer state sharded, reduce-scatter / all-gather around the AdamW kernel) makes the same
    update as the replicated optimizer behind the all-reduce, in eager steps and in one captured graph;
    each rank holds half of the moments."""
    base = _run("eager")
    for mode in ("zero_eager", "zero_graph"):
        res = _run(mode)
        assert res["equal"], f"parameters diverged across ranks by {res['maxdiff']} ({mode})"
        assert res["n_state"] * 2 <= res["n_params"] + 128
        diff = (torch.from_numpy(res["params"]) - torch.from_numpy(base["params"])).abs()
        # same mean gradient; only the clip coefficient's fp32 summation order differs
        assert float(diff.max()) <= 1e-2 * float(abs(base["params"]).max())
        assert float((diff > 0).float().mean()) < 0.2


# ======================================================================================
# Against the REFERENCE's own multi-GPU machinery (baseline/_ref + torch's DDP / ZeroRedundancyOptimizer
# exactly as build_components.py:176 and :250-256 construct them), 2 ranks over NCCL.
# ======================================================================================
TINY = dict(vocab_size=1024, context_length=256, emb_dim=256, n_heads=4, n_layers=2, hidden_dim=512,
            n_kv_groups=2, rope_base=10000.0, rope_freq=None, dtype=torch.bfloat16)


def _init(rank, world, port):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    return torch.device("cuda", rank)


def _rank_batch(rank, B=2, T=256, V=1024):
    g = torch.Generator().manual_seed(500 + rank)
    x = torch.randint(0, V, (B, T), generator=g)
    y = torch.randint(0, V, (B, T), generator=g)
    y[:, :60 + 90 * rank] = -100          # uneven masking: mean of per-rank means != global token mean
    return x, y


def _worker_vs_reference_ddp(rank, world, port, out):
    import contextlib
    import copy
    import io
    dev = _init(rank, world, port)
    from baseline import reference_arm as RA
    from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model, SharedBuffers
    from building_llm_from_scratch_b200.ddp import B200DDP
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    m = RA.modules()
    torch.manual_seed(7)                      # same init on both ranks
    m.llama3.SharedBuffers._buffers.clear()
    with contextlib.redirect_stdout(io.StringIO()):
        ref = m.llama3.Llama3Model(copy.deepcopy(TINY))
    for p in ref.parameters():
        p.requires_grad = False
    m.lora.replace_linear_with_lora(ref, rank=16, alpha=32, dtype=torch.bfloat16)
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if n.endswith("lora.B"):
                p.normal_(0, 0.02)
    state = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    ref.to(dev).train()
    from torch.nn.parallel import DistributedDataParallel as DDP
    ref_ddp = DDP(ref, device_ids=[rank])                                     # build_components.py:176
    x, y = _rank_batch(rank)
    logits = ref_ddp(x.to(dev))
    loss_ref = torch.nn.functional.cross_entropy(logits.flatten(0, 1), y.to(dev).flatten())      # train.py:91
    loss_ref.backward()
    ref_grads = {n: p.grad.detach().float().cpu() for n, p in ref.named_parameters() if p.grad is not None}

    SharedBuffers._buffers.clear()
    ours = Llama3Model(copy.deepcopy(TINY))
    for p in ours.parameters():
        p.requires_grad = False
    replace_linear_with_lora(ours, rank=16, alpha=32, dtype=torch.bfloat16)
    ours.load_state_dict(state)
    ours.to(dev).train()
    ddp = B200DDP(ours)
    opt = B200AdamW(ours.parameters(), lr=1e-3, weight_decay=0.1)
    opt.zero_grad()
    loss = ours.forward_loss(x.to(dev), y.to(dev))
    loss.backward()
    ddp.reduce_gradients(opt.flat_grads)
    torch.cuda.synchronize()
    worst, name = 0.0, None
    from conftest import rel_err
    for n, p in ours.named_parameters():
        if not p.requires_grad:
            continue
        e = rel_err(p.main_grad.float().cpu(), ref_grads[n])
        if e > worst:
            worst, name = e, n
    if rank == 0:
        out.put(dict(worst=worst, name=name, n=len(ref_grads), loss=float(loss), loss_ref=float(loss_ref)))
    dist.barrier()
    dist.destroy_process_group()


def _worker_zero_vs_reference(rank, world, port, out):
    dev = _init(rank, world, port)
    from torch.distributed.optim import ZeroRedundancyOptimizer
    from building_llm_from_scratch_b200.optim import B200AdamW
    g0 = torch.Generator().manual_seed(11)
    shapes = [(64, 16), (16, 64), (300,), (128, 128)]
    init = [(torch.randn(s, generator=g0) * 0.1).to(torch.bfloat16) for s in shapes]
    ref_p = [torch.nn.Parameter(t.clone().to(dev)) for t in init]
    our_p = [torch.nn.Parameter(t.clone().to(dev)) for t in init]
    zopt = ZeroRedundancyOptimizer(ref_p, optimizer_class=torch.optim.AdamW, lr=1e-2, weight_decay=0.1)   # build_components.py:250-256
    opt = B200AdamW(our_p, lr=1e-2, weight_decay=0.1, zero=True)
    opt.max_norm = 1.0
    gr = torch.Generator().manual_seed(100 + rank)
    for step in range(4):
        local = [(torch.randn(s, generator=gr) * (3.0 if step == 0 else 0.02)).to(torch.bfloat16).to(dev) for s in shapes]
        # reference: DDP leaves the bf16 mean over ranks in .grad, then clip 1.0 (train.py:115-120), then the sharded step
        for p, gl in zip(ref_p, local):
            avg = gl.clone()
            dist.all_reduce(avg, op=dist.ReduceOp.AVG)
            p.grad = avg
        torch.nn.utils.clip_grad_norm_(ref_p, max_norm=1.0)
        zopt.step()
        # ours: local gradients in main_grad; the optimizer reduce-scatters, clips and all-gathers itself
        opt.zero_grad()
        for p, gl in zip(our_p, local):
            p.main_grad.copy_(gl.float())
        opt.step()
    torch.cuda.synchronize()
    from conftest import rel_err
    errs = [rel_err(a.detach().float().cpu(), b.detach().float().cpu()) for a, b in zip(our_p, ref_p)]
    moved = [float((b.detach().float().cpu() - t.float()).abs().max()) for b, t in zip(ref_p, init)]
    flat = torch.cat([p.detach().float().reshape(-1) for p in our_p])
    both = [torch.zeros_like(flat) for _ in range(world)]
    dist.all_gather(both, flat)
    if rank == 0:
        out.put(dict(errs=errs, moved=moved, same=bool(torch.equal(both[0], both[1]))))
    dist.barrier()
    dist.destroy_process_group()


def _worker_zero_resume(rank, world, port, out):
    import tempfile
    dev = _init(rank, world, port)
    from building_llm_from_scratch_b200.Models.Llama.Llama3 import Llama3Model, SharedBuffers
    from building_llm_from_scratch_b200.ddp import B200DDP
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    tmp = [tempfile.mkdtemp() if rank == 0 else None]
    dist.broadcast_object_list(tmp, src=0)
    save_dir = tmp[0]

    def fresh():
        import copy
        torch.manual_seed(3)
        SharedBuffers._buffers.clear()
        model = Llama3Model(copy.deepcopy(TINY))
        for p in model.parameters():
            p.requires_grad = False
        replace_linear_with_lora(model, rank=16, alpha=32, dtype=torch.bfloat16)
        with torch.no_grad():
            for n, p in model.named_parameters():
                if n.endswith("lora.B"):
                    p.normal_(0, 0.02)
        model.to(dev).train()
        ddp = B200DDP(model)
        opt = B200AdamW(model.parameters(), lr=2e-3, weight_decay=0.1, zero=True)
        tr = Trainer(ddp, opt, TINY, [], None, save_dir, device=dev, warmup_steps=2, rank=rank)
        tr.set_schedule(total_training_steps=50, peak_lr=2e-3)
        return tr, opt

    batches = []
    g = torch.Generator().manual_seed(900 + rank)
    for _ in range(5):
        x = torch.randint(0, 1024, (2, 256), generator=g)
        y = torch.randint(0, 1024, (2, 256), generator=g)
        y[:, :50 * (rank + 1)] = -100
        batches.append((x.to(dev), y.to(dev)))
    tr, opt = fresh()                                  # uninterrupted run: 5 steps, checkpoint after 3
    start = torch.cat([g_["p"].float().reshape(-1) for g_ in opt._groups.values()]).clone()
    for i in range(3):
        tr.train_batch(*batches[i])
    tr.save_training_state("zero_ckpt.pt", asynchronous=False)
    dist.barrier()
    for i in range(3, 5):
        tr.train_batch(*batches[i])
    torch.cuda.synchronize()
    want = torch.cat([g_["p"].float().reshape(-1) for g_ in opt._groups.values()]).clone()
    want_m = torch.cat([g_["m"].float().reshape(-1) for g_ in opt._groups.values()]).clone()
    tr2, opt2 = fresh()                                # resumed run: load, then the last 2 steps
    tr2.load_training_state("zero_ckpt.pt")
    for i in range(3, 5):
        tr2.train_batch(*batches[i])
    torch.cuda.synchronize()
    got = torch.cat([g_["p"].float().reshape(-1) for g_ in opt2._groups.values()])
    got_m = torch.cat([g_["m"].float().reshape(-1) for g_ in opt2._groups.values()])
    ok = torch.tensor([int(torch.equal(want, got) and torch.equal(want_m, got_m))], device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)          # every rank resumed ITS shard of the moments
    if rank == 0:
        out.put(dict(equal=bool(ok.item()), maxdiff=float((want - got).abs().max()), step=opt2._step,
                     moved=float((want - start).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def _run_worker(fn):
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=fn, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    import time
    res, t0 = None, time.time()
    while res is None and time.time() - t0 < 200:
        if not out.empty():
            res = out.get()
        elif any(p.exitcode not in (None, 0) for p in procs):
            break
        else:
            time.sleep(0.2)
    for p in procs:
        p.join(timeout=60)
        if p.exitcode is None:
            p.kill()
        assert p.exitcode == 0, f"rank exited with {p.exitcode}"
    assert res is not None, "no result from rank 0"
    return res


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_b200ddp_gradients_match_reference_ddp():
    """The averaged LoRA gradients B200DDP leaves in main_grad against the reference model under torch's DDP
    (baseline/_ref, NCCL, uneven -100 masking across the two ranks): mean of per-rank means, <= 5e-2 (the
    suite's bound for bf16 gradients through a whole model)."""
    from baseline import reference_arm as RA
    if not RA.available():
        pytest.skip("baseline/_ref is not installed (run __graft_entry__.build() where the reference is mounted)")
    res = _run_worker(_worker_vs_reference_ddp)
    assert res["n"] > 0
    assert abs(res["loss"] - res["loss_ref"]) <= 2e-2 * abs(res["loss_ref"])
    assert res["worst"] <= 5e-2, f"{res['name']}: {res['worst']}"


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_zero1_matches_reference_zero_redundancy_optimizer():
    """B200AdamW(zero=True) against torch's ZeroRedundancyOptimizer(AdamW) built as build_components.py:250-256
    does, fed the same per-rank gradients (first step clipped, later ones not): parameters after 4 steps within
    the bf16 bound, identical on both ranks."""
    res = _run_worker(_worker_zero_vs_reference)
    assert res["same"]
    assert min(res["moved"]) > 1e-3                  # the steps did move every parameter
    assert max(res["errs"]) <= 2e-2, res["errs"]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_zero1_checkpoint_resume_is_exact():
    """save_training_state under ZeRO-1 gathers every rank's shard of the moments; a resumed run (each rank taking
    its own slice back) ends bit-identical to the uninterrupted one."""
    res = _run_worker(_worker_zero_resume)
    assert res["moved"] > 0
    assert res["equal"], f"resumed run differs by {res['maxdiff']}"
    assert res["step"] == 5
