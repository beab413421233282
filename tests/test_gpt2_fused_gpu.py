"""GPT-2 fusions (csrc/ops/transformer.cu, parallel/zero.py): residual add folded into LayerNorm, the one-kernel token +
position embedding with its scatter backward, and the bf16 weight shadows kept by the fused sharded optimizer - each against
plain PyTorch fp32."""
import os

import pytest
import torch
import torch.nn.functional as F

import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu]


@pytest.mark.parametrize("dtype,C", [(torch.float32, 256), (torch.bfloat16, 768), (torch.bfloat16, 1024)])
def test_add_layer_norm_matches_torch(dtype, C):
    from distributed_compute_pytorch_b200.ops import cuda as K
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(0)
    rows = 333
    x = torch.randn(rows, C, generator=g).to(dev, dtype).requires_grad_(True)
    r = torch.randn(rows, C, generator=g).to(dev, dtype).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(C, generator=g)).to(dev).requires_grad_(True)
    beta = (0.1 * torch.randn(C, generator=g)).to(dev).requires_grad_(True)
    ws, wy = torch.randn(rows, C, generator=g).to(dev), torch.randn(rows, C, generator=g).to(dev)
    assert K.add_layer_norm_supported(x)
    s, y = K.add_layer_norm(x, r, gamma, beta, 1e-5)
    assert s.dtype == dtype and y.dtype == dtype
    ((s.float() * ws).sum() + (y.float() * wy).sum()).backward()
    # fp32 reference on the same (rounded) inputs; the sum is rounded to the activation dtype before it is normalised
    xr, rr = x.detach().float().requires_grad_(True), r.detach().float().requires_grad_(True)
    gr, br = gamma.detach().clone().requires_grad_(True), beta.detach().clone().requires_grad_(True)
    sr = xr + rr
    yr = F.layer_norm(sr, (C,), gr, br, 1e-5)
    ((sr * ws).sum() + (yr * wy).sum()).backward()
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert torch.allclose(s.float(), sr, atol=tol, rtol=tol)
    assert torch.allclose(y.float(), yr, atol=4 * tol, rtol=tol)
    assert x.grad.data_ptr() == r.grad.data_ptr() or torch.equal(x.grad, r.grad)
    assert torch.allclose(x.grad.float(), xr.grad, atol=8 * tol, rtol=4 * tol), (x.grad.float() - xr.grad).abs().max().item()
    assert torch.allclose(gamma.grad, gr.grad, atol=rows * tol * 0.2, rtol=4 * tol)
    assert torch.allclose(beta.grad, br.grad, atol=rows * tol * 0.2, rtol=4 * tol)


def test_add_layer_norm_unused_outputs():
    """Only one of the two outputs may carry a gradient (the last LayerNorm of the model drops the sum)."""
    from distributed_compute_pytorch_b200.ops import cuda as K
    dev = torch.device("cuda", 0)
    x = torch.randn(64, 256, device=dev, requires_grad=True); r = torch.randn(64, 256, device=dev, requires_grad=True)
    gamma = torch.ones(256, device=dev, requires_grad=True); beta = torch.zeros(256, device=dev, requires_grad=True)
    _, y = K.add_layer_norm(x, r, gamma, beta)
    y.square().sum().backward()
    xr = (x.detach() + r.detach()).requires_grad_(True)
    F.layer_norm(xr, (256,)).square().sum().backward()
    assert torch.allclose(x.grad, xr.grad, atol=1e-4, rtol=1e-4)
    x.grad = None; r.grad = None
    s, _ = K.add_layer_norm(x, r, gamma, beta)
    s.sum().backward()
    assert torch.equal(x.grad, torch.ones_like(x)) and torch.equal(r.grad, torch.ones_like(r))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_embedding_matches_torch(dtype):
    from distributed_compute_pytorch_b200.ops import cuda as K
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(1)
    V, P, C, B, T = 1000, 64, 96, 3, 40
    wte = torch.randn(V, C, generator=g).to(dev).requires_grad_(True)
    wpe = torch.randn(P, C, generator=g).to(dev).requires_grad_(True)
    idx = torch.randint(0, 50, (B, T), generator=g).to(dev)            # few distinct tokens: many collisions in the scatter
    w = torch.randn(B, T, C, generator=g).to(dev)
    out = K.embed(idx, wte, wpe, dtype)
    assert out.dtype == dtype and out.shape == (B, T, C)
    (out.float() * w).sum().backward()
    wr, pr = wte.detach().clone().requires_grad_(True), wpe.detach().clone().requires_grad_(True)
    ref = F.embedding(idx, wr) + F.embedding(torch.arange(T, device=dev), pr)
    (ref.to(dtype).float() * w).sum().backward()
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert torch.allclose(out.float(), ref, atol=tol, rtol=tol)
    gtol = 1e-4 if dtype == torch.float32 else 5e-2
    assert torch.allclose(wte.grad, wr.grad, atol=gtol, rtol=gtol), (wte.grad - wr.grad).abs().max().item()
    assert torch.allclose(wpe.grad, pr.grad, atol=gtol, rtol=gtol)
    assert wpe.grad[T:].abs().max().item() == 0.0 and wte.grad[50:].abs().max().item() == 0.0


def test_gpt2_fused_path_matches_unfused():
    """The model with embedding / add+LayerNorm fusions against the same weights through the layer-by-layer path."""
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    cfg = dcp.models.GPT2Config(vocab_size=512, n_positions=128, n_embd=128, n_layer=2, n_head=2)
    model = dcp.models.GPT2(cfg).to(dev)
    model.compute_dtype = torch.bfloat16
    idx = torch.randint(0, 512, (2, 128), device=dev); tgt = torch.randint(0, 512, (2, 128), device=dev)
    res = {}
    for fused in ("1", "0"):
        os.environ["DCP_GPT2_FUSED"] = fused
        try:
            model.zero_grad(set_to_none=True)
            loss = model(idx, tgt)
            loss.backward()
            res[fused] = (loss.item(), {n: p.grad.clone() for n, p in model.named_parameters()})
        finally:
            os.environ.pop("DCP_GPT2_FUSED", None)
    assert abs(res["1"][0] - res["0"][0]) < 2e-2, (res["1"][0], res["0"][0])
    for n, g1 in res["1"][1].items():
        g0 = res["0"][1][n]
        den = g0.abs().max().item() + 1e-6
        assert (g1 - g0).abs().max().item() / den < 6e-2, (n, (g1 - g0).abs().max().item(), den)


def _shadow_worker(rank, world, port):
    from distributed_compute_pytorch_b200.ops import cuda as K
    from distributed_compute_pytorch_b200.parallel.zero import FusedShardedOptimizer
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    model = torch.nn.Sequential(dcp.ops.Linear(64, 256), dcp.ops.ReLU(), dcp.ops.Linear(256, 16)).to(dev)
    ddp = dcp.DistributedDataParallel(model)
    opt = FusedShardedOptimizer(ddp, kind="adamw", lr=1e-2, shadow_dtype=torch.bfloat16)
    w = model[0].weight
    assert (w.data_ptr(), torch.bfloat16) in K._SHADOWS
    x = torch.randn(32, 64, device=dev, dtype=torch.bfloat16); y = torch.randint(0, 16, (32,), device=dev)
    for _ in range(3):
        opt.zero_grad()
        before = w.detach().clone()
        F.cross_entropy(ddp(x).float(), y).backward()
        opt.step()
        torch.cuda.synchronize()
        assert not torch.equal(before, w.detach())                                   # the update ran
        sh = K._SHADOWS[(w.data_ptr(), torch.bfloat16)][0]
        assert torch.equal(sh, w.detach().to(torch.bfloat16))                          # ... and the shadow followed it
        assert K._lowp(w.data, torch.bfloat16).data_ptr() == sh.data_ptr()             # which is what the Linear multiplies with
    # weights written through load_state_dict reach the shadow as well
    sd = {k: torch.full_like(v, 0.25) for k, v in model.state_dict().items()}
    model.load_state_dict(sd)
    torch.cuda.synchronize()
    assert torch.equal(K._SHADOWS[(w.data_ptr(), torch.bfloat16)][0], torch.full_like(w, 0.25, dtype=torch.bfloat16))
    opt.close()
    assert (w.data_ptr(), torch.bfloat16) not in K._SHADOWS
    dist.destroy_process_group()


def test_bf16_shadow_follows_the_fused_optimizer():
    dcp.spawn(_shadow_worker, args=(1, free_port()), nprocs=1, join=True)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gelu_backward_hands_the_bias_gradient_to_the_linear(dtype):
    """GELU backward sums its dx over the rows; the Linear below takes that as its bias gradient (ops/cuda.py _COLSUMS)."""
    from distributed_compute_pytorch_b200.ops import cuda as K
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    rows, Cin, Cout = 1000, 64, 256                      # 1000 rows: ragged against the 16-row strips of the kernel
    lin = dcp.ops.Linear(Cin, Cout).to(dev)
    act = dcp.ops.GELU()
    x = torch.randn(rows, Cin, device=dev, dtype=dtype)
    w = torch.randn(rows, Cout, device=dev)
    K._COLSUMS.clear()
    (act(lin(x)).float() * w).sum().backward()
    assert not K._COLSUMS                                  # the offer was taken
    ref = torch.nn.Linear(Cin, Cout).to(dev)
    ref.load_state_dict(lin.state_dict())
    (F.gelu(ref(x.float()), approximate="tanh") * w).sum().backward()
    tol = 2e-3 if dtype == torch.float32 else 6e-2
    den = ref.bias.grad.abs().max().item()
    assert (lin.bias.grad - ref.bias.grad).abs().max().item() / den < tol
    assert (lin.weight.grad - ref.weight.grad).abs().max().item() / ref.weight.grad.abs().max().item() < tol
