"""GPU: the fp32 tolerance class of the north star — per-op outputs and gradients within 1e-4 relative — on the
fixtures recorded from the reference's modules in float32 (GPT-2 is fp32 as written, Models/GPT2/GPT2.py:20-23; the
Llama sub-modules were built with dtype=float32). The drop-in modules are fed float32 tensors, which routes every op
to the plain-fp32 kernels of csrc/fp32.cu (FFMA GEMM, materialised-score attention, LayerNorm, GELU, SwiGLU, RoPE,
cross-entropy, embedding) through the same C-ABI library."""
import copy

import pytest
import torch

from conftest import close, rel_err

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]
FP32_TOL = 1e-4


def _check_module(mod, f, tol=FP32_TOL):
    x = f["x"].cuda().requires_grad_(True)
    y = mod(x)
    assert y.dtype == torch.float32
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu(), f["y"]) < tol
    assert rel_err(x.grad.cpu(), f["dx"]) < tol
    params = dict(mod.named_parameters())
    for k, v in f["grads"].items():
        assert close(params[k].grad.cpu(), v, tol, atol=1e-7), (k, rel_err(params[k].grad.cpu(), v))


def test_gemm_f32_strided_and_batched_against_fp64():
    from building_llm_from_scratch_b200 import f32
    g = torch.Generator(device="cuda").manual_seed(0)
    a = torch.randn(300, 515, generator=g, device="cuda")
    b = torch.randn(190, 515, generator=g, device="cuda")
    bias = torch.randn(190, generator=g, device="cuda")
    ref = (a.double() @ b.double().t() + bias.double()).float()
    assert rel_err(f32.gemm(a, b, bias), ref) < 1e-6
    at, bt = a.t().contiguous().t(), b.t().contiguous().t()          # column-major views of the same matrices
    assert rel_err(f32.gemm(at, bt, bias), ref) < 1e-6
    out = torch.ones(300, 190, device="cuda")
    f32.gemm(a, b, out=out, alpha=0.5, beta=2.0)
    assert rel_err(out, (0.5 * (a.double() @ b.double().t()) + 2.0).float()) < 1e-6
    A3, B3 = torch.randn(5, 70, 33, generator=g, device="cuda"), torch.randn(1, 41, 33, generator=g, device="cuda").expand(5, 41, 33)
    O3 = torch.empty(5, 70, 41, device="cuda")
    f32.bgemm(A3, B3, O3)
    assert rel_err(O3, torch.einsum("bmk,bnk->bmn", A3.double(), B3.double()).float()) < 1e-6


def test_gpt2_modules_fp32_against_reference_fixtures(golden):
    from building_llm_from_scratch_b200.Models.GPT2 import GPT2 as G
    g = golden("gpt2.pt")
    f = g["mha/fp32"]
    mha = G.MultiHeadAttention(128, 128, 64, 0.0, f["n_heads"], qkv_bias=True).cuda()
    mha.load_state_dict(f["sd"])
    _check_module(mha, f)
    f = g["ffn/fp32"]
    ff = G.FeedForward(dict(emb_dim=64)).cuda()
    ff.load_state_dict(f["sd"])
    _check_module(ff, f)


@pytest.mark.parametrize("which", ["model", "model_lora"])
def test_gpt2_model_fp32_against_reference_fixture(golden, which):
    """Whole GPT-2 (Models/GPT2/GPT2.py:91-124) in fp32, full training and LoRA: logits, loss and every gradient at 1e-4."""
    from building_llm_from_scratch_b200.Models.GPT2 import GPT2 as G
    from building_llm_from_scratch_b200.lora import replace_linear_with_lora
    f = golden("gpt2.pt")[which]
    model = G.GPTModel(copy.deepcopy(f["cfg"]))
    if which == "model_lora":
        for p in model.parameters():
            p.requires_grad = False
        replace_linear_with_lora(model, rank=f["rank"], alpha=f["alpha"], dtype=torch.float32)
    model.load_state_dict(f["sd"])
    model.cuda().train()
    x, y = f["x"].cuda(), f["y"].cuda()
    logits = model(x)
    assert logits.dtype == torch.float32
    assert rel_err(logits.cpu(), f["logits"]) < FP32_TOL
    loss = model.forward_loss(x, y)
    assert abs(float(loss) - float(f["loss"])) <= FP32_TOL * abs(float(f["loss"]))
    loss.backward()
    params = dict(model.named_parameters())
    assert set(f["grads"]) == {k for k, p in params.items() if p.requires_grad}
    for k, v in f["grads"].items():
        assert close(params[k].grad.cpu(), v, FP32_TOL, atol=1e-8), (k, rel_err(params[k].grad.cpu(), v))


def test_llama_modules_fp32_against_reference_fixtures(golden):
    from building_llm_from_scratch_b200.Models.Llama import Llama3 as L
    from building_llm_from_scratch_b200.Models.Llama.common_components import FeedForward, RMSNorm, compute_rope
    g = golden("llama_modules.pt")
    f = g["rmsnorm/torch.float32"]
    n = RMSNorm(64).cuda()
    n.load_state_dict({"weight": f["w"]})
    x = f["x"].cuda().requires_grad_(True)
    y = n(x)
    y.backward(f["gy"].cuda())
    assert rel_err(y.cpu(), f["y"]) < FP32_TOL and rel_err(x.grad.cpu(), f["dx"]) < FP32_TOL
    assert rel_err(n.weight.grad.cpu(), f["dw"]) < FP32_TOL
    f = g["rope/torch.float32"]
    assert rel_err(compute_rope(f["x"].cuda(), f["cos"].cuda(), f["sin"].cuda()).cpu(), f["y"]) < 1e-6
    f = g["ffn/torch.float32"]
    ff = FeedForward(dict(emb_dim=64, hidden_dim=160, dtype=torch.float32)).cuda()
    ff.load_state_dict(f["sd"])
    _check_module(ff, f)
    f = g["gqa/torch.float32"]
    L.SharedBuffers._buffers.clear()
    fc = dict(factor=32.0, low_freq_factor=1.0, high_freq_factor=4.0, original_context_length=8192)
    att = L.GroupedQueryAttention(256, 256, 64, f["n_heads"], f["n_kv_groups"], rope_base=31250.0, rope_config=fc,
                                  dtype=torch.float32).cuda()
    att.load_state_dict(f["sd"])
    _check_module(att, f)


def test_gpt2_fp32_training_steps_with_fused_optimizer(golden):
    """A few Trainer.train_batch steps of the fp32 GPT-2 fixture model through B200AdamW (fp32 parameters, main_grad
    accumulation by the fp32 GEMM) against the same steps with torch.optim.AdamW + clip_grad_norm_ on autograd's
    gradients: the losses agree to 1e-4."""
    from building_llm_from_scratch_b200.Models.GPT2 import GPT2 as G
    from building_llm_from_scratch_b200.optim import B200AdamW
    from building_llm_from_scratch_b200.train import Trainer
    f = golden("gpt2.pt")["model"]
    x, y = f["x"].cuda(), f["y"].cuda()

    def run(fused):
        model = G.GPTModel(copy.deepcopy(f["cfg"]))
        model.load_state_dict(f["sd"])
        model.cuda().train()
        opt = (B200AdamW if fused else torch.optim.AdamW)(model.parameters(), lr=1e-3, weight_decay=0.1)
        tr = Trainer(model, opt, f["cfg"], [], None, "/tmp", device="cuda", warmup_steps=2)
        tr.set_schedule(20, peak_lr=1e-3)
        return [float(tr.train_batch(x, y)) for _ in range(4)]
    a, b = run(True), run(False)
    assert a[0] > a[-1]
    for u, v in zip(a, b):
        assert abs(u - v) <= 2e-4 * abs(v), (a, b)
