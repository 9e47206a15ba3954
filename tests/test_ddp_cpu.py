"""CPU, world_size 2 over gloo: the data-parallel exchange of ddp.B200DDP reproduces the reference
DDP's semantics: every rank ends with the UNWEIGHTED MEAN over ranks of the per-rank
gradients (each already a local mean over that rank's non-ignored targets), which differs from the
global token mean when the ranks have different numbers of valid targets."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from building_llm_from_scratch_b200.ddp import B200DDP
    from oracle import reference_ops as R
    torch.manual_seed(0)
    lin = torch.nn.Linear(16, 32, bias=False)
    if rank == 1:                               # ranks start different; the wrapper broadcasts rank 0's
        with torch.no_grad():
            lin.weight.add_(1.0)
    model = B200DDP(lin, bucket_mb=0.001)       # tiny buckets -> several all-reduces
    g = torch.Generator().manual_seed(100 + rank)
    x = torch.randn(8, 16, generator=g)
    y = torch.randint(0, 32, (8,), generator=g)
    if rank == 1:
        y[:6] = -100                            # uneven masking across ranks
    loss = R.cross_entropy(model(x), y)
    loss.backward()
    local = lin.weight.grad.clone()
    flat = lin.weight.grad.reshape(-1).clone().float()
    model.reduce_gradients([flat])
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    weights = [torch.zeros_like(lin.weight) for _ in range(world)]
    dist.all_gather(weights, lin.weight.detach())
    if rank == 0:
        out.put(dict(reduced=flat.view_as(local), locals=gathered, weights=weights))
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_mean_of_means_two_ranks():
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get()
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert torch.equal(res["weights"][0], res["weights"][1]), "parameters must be broadcast from rank 0"
    expect = (res["locals"][0] + res["locals"][1]) / 2
    assert torch.allclose(res["reduced"], expect, rtol=1e-6, atol=1e-7)
    # and it is NOT the valid-token-weighted mean (8 valid rows on rank 0, 2 on rank 1)
    weighted = (8 * res["locals"][0] + 2 * res["locals"][1]) / 10
    assert not torch.allclose(res["reduced"], weighted, rtol=1e-3, atol=1e-5)


# ---- overlapped, bucketed exchange (attach / begin_backward / finish_backward) -------------------
class _FakeOpt:
    """The part of optim.B200AdamW's layout that B200DDP.attach() reads."""
    def __init__(self, params):
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += (p.numel() + 63) // 64 * 64
        flat = torch.zeros(total)
        for p, o in zip(params, offs):
            p.main_grad = flat[o:o + p.numel()].view(p.shape)
        self._groups = {torch.float32: dict(params=params, offsets=offs, g=flat)}

    def flush_pending_grads(self):
        pass


def _overlap_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from building_llm_from_scratch_b200.ddp import B200DDP
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(16, 24), torch.nn.Linear(24, 40), torch.nn.Linear(40, 8))
    params = list(net.parameters())
    model = B200DDP(net, bucket_mb=0.004)        # ~1000 floats per bucket -> several buckets
    opt = _FakeOpt(params)
    model.attach(opt)
    n_buckets = len(model._buckets)
    sizes = [b["hi"] - b["lo"] for b in model._buckets]
    cover = sorted((b["lo"], b["hi"]) for b in model._buckets)
    # step: each rank's "backward" writes rank-dependent gradients and reports parameters in reverse order,
    # the last Linear's bias through the autograd-hook route, one parameter not at all (unused)
    model.begin_backward()
    issued_during = []
    for i, p in enumerate(reversed(params)):
        if i == len(params) - 1:
            break                                  # first parameter never reports: finish_backward sends it
        g = torch.full(p.shape, float(rank + 1) * (i + 1))
        if i == 0:
            p.grad = g.clone()
            model._autograd_hook(p)
            assert p.grad is None
        else:
            p.main_grad.copy_(g)
            model._on_ready([p])
        issued_during.append(sum(b["issued"] for b in model._buckets))
    try:
        model._on_ready([params[-1]])
        dup = False
    except RuntimeError:
        dup = True
    model.finish_backward()
    flat = opt._groups[torch.float32]["g"]
    got = [p.main_grad.clone() for p in params]
    if rank == 0:
        out.put(dict(n_buckets=n_buckets, sizes=sizes, cover=cover, total=flat.numel(), issued=issued_during,
                     got=got, dup=dup))
    dist.barrier()
    dist.destroy_process_group()


def test_overlapped_buckets_two_ranks():
    ctx = mp.get_context("spawn")
    out = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_overlap_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get()
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res["n_buckets"] >= 2
    # buckets tile the flat buffer exactly
    pos = 0
    for lo, hi in res["cover"]:
        assert lo == pos
        pos = hi
    assert pos == res["total"]
    # all-reduces were issued while "backward" was still reporting, not only at the end
    assert res["issued"][-1] >= 1 and res["issued"][0] <= res["issued"][-1]
    assert res["dup"], "a second report of the same parameter must be refused"
    # mean over the two ranks of (rank + 1) * (i + 1) = 1.5 * (i + 1); the unreported first parameter stays 0
    n = len(res["got"])
    for i, g in enumerate(reversed(res["got"])):
        expect = 0.0 if i == n - 1 else 1.5 * (i + 1)
        assert torch.allclose(g, torch.full_like(g, expect)), (i, g.flatten()[:4])
