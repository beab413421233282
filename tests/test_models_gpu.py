"""North-star model families on the hand-written kernels: ResNet-18 / ResNet-50 (BatchNorm, residuals, strided and
1x1 convs, bf16 compute with fp32 master weights) and GPT-2 (LayerNorm, GELU, tensor-core Linear), each against a
pure-PyTorch CPU fp32 run of the same module, plus a short training sanity check."""
import copy

import pytest
import torch
import torch.nn.functional as F

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def test_resnet18_cifar_forward_backward_matches_cpu():
    torch.manual_seed(0)
    cpu = dcp.models.resnet18(num_classes=10).train()
    gpu = copy.deepcopy(cpu).to(DEV).train()
    x = torch.randn(16, 3, 32, 32); y = torch.randint(0, 10, (16,))
    lc = F.cross_entropy(cpu(x), y); lc.backward()
    lg = ops.cross_entropy(gpu(x.to(DEV)), y.to(DEV)); lg.backward()
    assert abs(lg.item() - lc.item()) < 2e-2 * max(1.0, abs(lc.item()))
    errs = {n: _rel(p.grad, q.grad) for (n, p), q in zip(gpu.named_parameters(), cpu.parameters())}
    bad = {n: e for n, e in errs.items() if e > 0.15}
    assert not bad, bad                                    # tf32 + ReLU-mask flips through 20 layers
    assert _rel(gpu.bn1.running_mean, cpu.bn1.running_mean) < 1e-3
    sd = gpu.state_dict()
    assert "layer2.0.downsample.0.weight" in sd and tuple(sd["fc.weight"].shape) == (10, 512)


def test_resnet50_bf16_compute_runs_and_tracks_fp32():
    torch.manual_seed(0)
    m32 = dcp.models.resnet50(num_classes=100).to(DEV).train()
    m16 = copy.deepcopy(m32); m16.compute_dtype = torch.bfloat16
    x = torch.randn(4, 3, 64, 64, device=DEV); y = torch.randint(0, 100, (4,), device=DEV)
    l32 = ops.cross_entropy(m32(x), y); l32.backward()
    l16 = ops.cross_entropy(m16(x), y); l16.backward()
    assert torch.isfinite(l16) and abs(l16.item() - l32.item()) < 0.1 * max(1.0, abs(l32.item()))
    assert m16.conv1.weight.grad.dtype == torch.float32 and m16.fc.weight.grad.dtype == torch.float32   # fp32 master grads
    assert torch.isfinite(m16.fc.weight.grad).all() and _rel(m16.fc.weight.grad, m32.fc.weight.grad) < 0.6   # bf16 through 53 convs, batch-4 BN
    assert sum(p.numel() for p in m32.parameters()) == 25_557_032 - 2048 * 900 - 900


def test_gpt2_small_config_matches_cpu_and_bf16():
    torch.manual_seed(0)
    cfg = dcp.models.GPT2Config(vocab_size=1000, n_positions=64, n_embd=128, n_layer=2, n_head=4)
    cpu = dcp.models.GPT2(cfg)
    gpu = copy.deepcopy(cpu).to(DEV)
    idx = torch.randint(0, 1000, (4, 64)); tgt = torch.randint(0, 1000, (4, 64))
    lc = cpu(idx, tgt); lc.backward()
    lg = gpu(idx.to(DEV), tgt.to(DEV)); lg.backward()
    assert abs(lg.item() - lc.item()) < 5e-3 * abs(lc.item())
    for (n, p), q in zip(gpu.named_parameters(), cpu.parameters()):
        assert _rel(p.grad, q.grad) < 2e-2, (n, _rel(p.grad, q.grad))
    g16 = copy.deepcopy(cpu).to(DEV); g16.compute_dtype = torch.bfloat16
    l16 = g16(idx.to(DEV), tgt.to(DEV)); l16.backward()
    assert abs(l16.item() - lc.item()) < 3e-2 * abs(lc.item())
    assert g16.h[0].attn_qkv.weight.grad.dtype == torch.float32
    assert sum(p.numel() for p in dcp.models.GPT2().parameters()) == 124_439_808 + (50304 - 50257) * 768


@pytest.mark.parametrize("rows,C", [(64, 768), (4100, 256), (2500, 1024)])
def test_layernorm_gelu_kernels(rows, C):
    """LayerNorm forward / backward (warp-per-row 16-byte kernels where the row fits the registers, CTA-per-row otherwise) and
    tanh-GELU, f32 and bf16, against PyTorch on the CPU."""
    torch.manual_seed(0)
    for dtype, tol in ((torch.float32, 1e-5), (torch.bfloat16, 2e-2)):
        x = torch.randn(rows, C, requires_grad=True); g = torch.randn(rows, C)
        ln = ops.LayerNorm(C); ln.weight.data.uniform_(0.5, 1.5); ln.bias.data.uniform_(-0.5, 0.5)
        y = ln(x); y.backward(g)
        lg = copy.deepcopy(ln).to(DEV); xg = x.detach().to(DEV).to(dtype).requires_grad_(True)
        yg = lg(xg); yg.backward(g.to(DEV).to(dtype))
        assert _rel(yg, y) < max(tol, 1e-5) and _rel(xg.grad, x.grad) < max(5 * tol, 1e-4)
        assert _rel(lg.weight.grad, ln.weight.grad) < max(5 * tol, 1e-4) and _rel(lg.bias.grad, ln.bias.grad) < max(5 * tol, 1e-4)
        x2 = torch.randn(1000, requires_grad=True); y2 = F.gelu(x2, approximate="tanh"); y2.backward(torch.ones(1000))
        xg2 = x2.detach().to(DEV).to(dtype).requires_grad_(True); yg2 = ops.GELU()(xg2); yg2.backward(torch.ones(1000, device=DEV, dtype=dtype))
        assert _rel(yg2, y2) < max(tol, 1e-5) and _rel(xg2.grad, x2.grad) < max(tol, 1e-5)


def test_resnet18_trains_with_graphed_step():
    from distributed_compute_pytorch_b200.parallel.engine import GraphedTrainStep
    torch.manual_seed(0)
    m = dcp.models.resnet18(num_classes=10).to(DEV).train()
    opt = ops.SGD(m.parameters(), lr=0.05, momentum=0.9)
    ds = dcp.utils.SyntheticImages(512, (3, 32, 32), 10, seed=1)
    xs, ys = ds.images.to(DEV), ds.labels.to(DEV)
    step = GraphedTrainStep(m, opt, lambda o, t: ops.cross_entropy(o, t), xs[:64], ys[:64])
    losses = [step(xs[i:i + 64], ys[i:i + 64]).item() for _ in range(4) for i in range(0, 512, 64)]
    assert losses[-1] == losses[-1] and sum(losses[-4:]) < 0.8 * sum(losses[:4]), (losses[:4], losses[-4:])
