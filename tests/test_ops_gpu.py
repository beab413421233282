"""Every sm_100a compute kernel vs a plain PyTorch fp32 reference of the same op (CPU oracle),
forward and backward, over the reference model's exact shapes plus ResNet-style ones."""
import pytest
import torch
import torch.nn.functional as F

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _cmp(a, b, rtol=1e-4, atol=1e-4, what=""):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = (a - b).abs().max().item()
    ref = b.abs().max().item() + 1e-12
    assert torch.allclose(a, b, rtol=rtol, atol=atol * max(1.0, ref)), f"{what}: max abs err {err:.3e} (ref max {ref:.3e})"


def _cmp_norm(a, b, tol, what=""):
    """Relative Frobenius error - for tf32-multiplied results, where a ReLU mask may flip on a few elements
    whose pre-activation is ~0 (the same happens between cuDNN-tf32 and an fp32 reference)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = ((a - b).norm() / (b.norm() + 1e-30)).item()
    assert err < tol, f"{what}: relative Frobenius error {err:.3e} >= {tol}"


def _native_loaded():
    with open("/proc/self/maps") as f:
        return "libdcp_cuda.so" in f.read()


def test_native_library_is_what_runs():
    from distributed_compute_pytorch_b200.ops import cuda
    before = cuda.launch_count
    ops.log_softmax(torch.randn(4, 10, device=DEV))
    assert cuda.launch_count == before + 1 and _native_loaded()


@pytest.mark.parametrize("N,C,H,W,K,R,stride,pad,bias,relu", [
    (128, 1, 28, 28, 32, 3, 1, 0, True, True),      # conv1 of the reference (ref:main.py:23,32-33)
    (128, 32, 26, 26, 64, 3, 1, 0, True, True),     # conv2 of the reference (ref:main.py:24,34-35)
    (8, 3, 32, 32, 64, 7, 2, 3, False, False),      # ResNet stem on CIFAR shape
    (4, 64, 8, 8, 128, 3, 2, 1, False, False),      # 3x3 stride 2
    (4, 64, 8, 8, 128, 1, 2, 0, False, False),      # 1x1 stride-2 downsample
    (2, 48, 9, 7, 40, 3, 1, 1, True, False),        # odd sizes
])
def test_conv2d(N, C, H, W, K, R, stride, pad, bias, relu):
    torch.manual_seed(0)
    conv = ops.Conv2d(C, K, R, stride, pad, bias=bias, fuse_relu=relu)
    x = torch.randn(N, C, H, W, requires_grad=True)
    y_ref = conv(x)
    g = torch.randn_like(y_ref)
    y_ref.backward(g)
    ref = (y_ref, x.grad, conv.weight.grad, conv.bias.grad if bias else None)
    conv_g = ops.Conv2d(C, K, R, stride, pad, bias=bias, fuse_relu=relu).to(DEV)
    conv_g.load_state_dict(conv.state_dict())
    xg = x.detach().to(DEV).requires_grad_(True)
    yg = conv_g(xg)
    yg.backward(g.to(DEV))
    # tf32-class tolerance (the tensor-core path multiplies in tf32 like cuDNN's default).  With a fused ReLU the
    # backward also sees a few flipped masks where the tf32 pre-activation changes sign, hence the looser bound.
    tol = 5e-3
    btol = 6e-2 if relu else tol
    _cmp_norm(yg, ref[0], tol, "y"); _cmp_norm(xg.grad, ref[1], btol, "dx"); _cmp_norm(conv_g.weight.grad, ref[2], btol, "dw")
    if bias:
        _cmp_norm(conv_g.bias.grad, ref[3], btol, "db")


@pytest.mark.parametrize("M,K,N,relu", [(128, 9216, 128, False), (128, 128, 10, False), (64, 512, 1000, True), (33, 77, 19, False), (1024, 768, 2304, False)])
def test_linear(M, K, N, relu):
    torch.manual_seed(0)
    lin = ops.Linear(K, N, fuse_relu=relu)
    x = torch.randn(M, K, requires_grad=True)
    y = lin(x); g = torch.randn_like(y); y.backward(g)
    lg = ops.Linear(K, N, fuse_relu=relu).to(DEV); lg.load_state_dict(lin.state_dict())
    xg = x.detach().to(DEV).requires_grad_(True)
    yg = lg(xg); yg.backward(g.to(DEV))
    tol = 5e-3
    btol = 6e-2 if relu else tol
    _cmp_norm(yg, y, tol, "y"); _cmp_norm(xg.grad, x.grad, btol, "dx")
    _cmp_norm(lg.weight.grad, lin.weight.grad, btol, "dw"); _cmp_norm(lg.bias.grad, lin.bias.grad, btol, "db")


@pytest.mark.parametrize("shape,relu", [((128, 128), True), ((128, 128), False), ((16, 64, 12, 12), True), ((4, 200, 5, 7), False),
                                        ((8, 256, 14, 14), True), ((33, 72, 9, 9), False), ((2, 2048, 7, 7), True)])
def test_batchnorm_train_and_eval(shape, relu):
    torch.manual_seed(0)
    C = shape[1]
    cls = ops.BatchNorm1d if len(shape) == 2 else ops.BatchNorm2d
    bn = cls(C, fuse_relu=relu)
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.5, 0.5)
    bg = cls(C, fuse_relu=relu).to(DEV); bg.load_state_dict(bn.state_dict())
    x = torch.randn(*shape) * 2 + 1
    x.requires_grad_(True)
    for step in range(2):
        y = bn(x); g = torch.randn_like(y); x.grad = None; bn.zero_grad(); y.backward(g)
        xg = x.detach().to(DEV).requires_grad_(True)
        if len(shape) == 4:
            xg = xg.detach().contiguous(memory_format=torch.channels_last).requires_grad_(True)
        bg.zero_grad(); yg = bg(xg); yg.backward(g.to(DEV))
        _cmp(yg, y, what="y"); _cmp(xg.grad, x.grad, what="dx", atol=2e-4)
        _cmp(bg.weight.grad, bn.weight.grad, what="dgamma", atol=2e-4); _cmp(bg.bias.grad, bn.bias.grad, what="dbeta", atol=2e-4)
        _cmp(bg.running_mean, bn.running_mean, what="running_mean"); _cmp(bg.running_var, bn.running_var, what="running_var")
    assert bg.num_batches_tracked.item() == 2
    bn.eval(); bg.eval()
    _cmp(bg(x.detach().to(DEV)), bn(x.detach()), what="eval y")


def test_loss_kernels():
    torch.manual_seed(0)
    for M, C in ((128, 10), (64, 1000), (8, 50257), (16, 50304), (6, 4096)):     # the last two take the 16-byte single-pass kernels
        x = torch.randn(M, C, requires_grad=True); t = torch.randint(0, C, (M,))
        xg = x.detach().to(DEV).requires_grad_(True); tg = t.to(DEV)
        for red in ("mean", "sum"):
            x.grad = None; xg.grad = None
            l = F.nll_loss(F.log_softmax(x, 1), t, reduction=red); l.backward()
            lg = ops.nll_loss(ops.log_softmax(xg, 1), tg, reduction=red); lg.backward()
            _cmp(lg, l, what=f"nll {red}"); _cmp(xg.grad, x.grad, what=f"dlogits {red}", atol=1e-5)
            xg.grad = None
            lc = ops.cross_entropy(xg, tg, reduction=red); lc.backward()
            _cmp(lc, l, what=f"ce {red}"); _cmp(xg.grad, x.grad, what=f"ce grad {red}", atol=1e-5)
        assert ops.correct_count(xg, tg).item() == x.argmax(1).eq(t).sum().item()
    # bf16 logits over a padded 50k vocabulary (the GPT-2 head): vectorised single-pass kernels vs fp32 math on the same bf16 values
    xb = (torch.randn(32, 50304) * 3).to(torch.bfloat16); tb = torch.randint(0, 50257, (32,))
    xr = xb.float().requires_grad_(True); lr = F.cross_entropy(xr, tb); lr.backward()
    xg = xb.to(DEV).requires_grad_(True); lg = ops.cross_entropy(xg, tb.to(DEV)); lg.backward()
    assert abs(lg.item() - lr.item()) < 2e-3 * abs(lr.item())
    assert ((xg.grad.float().cpu() - xr.grad).norm() / xr.grad.norm()).item() < 1e-2


def test_pool_relu_dropout_gap():
    torch.manual_seed(0)
    x = torch.randn(8, 64, 24, 24, requires_grad=True)
    for k, s, p, nchw in ((2, 2, 0, True), (2, 2, 0, False), (3, 2, 1, False)):
        y = F.max_pool2d(x, k, s, p); g = torch.randn_like(y); x.grad = None; y.backward(g)
        xg = x.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        yg = ops.MaxPool2d(k, s, p, out_nchw=nchw)(xg); yg.backward(g.to(DEV))
        assert yg.is_contiguous() == nchw or yg.is_contiguous(memory_format=torch.channels_last)
        _cmp(yg, y, what="pool y"); _cmp(xg.grad, x.grad, what="pool dx")
    xg = x.detach().to(DEV).requires_grad_(True)
    y = F.relu(x); x.grad = None; g = torch.randn_like(y); y.backward(g)
    yg = ops.ReLU()(xg); yg.backward(g.to(DEV)); _cmp(yg, y); _cmp(xg.grad, x.grad)
    from distributed_compute_pytorch_b200.ops import cuda
    xg = x.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = x.mean((2, 3)); x.grad = None; g = torch.randn_like(y); y.backward(g)
    yg = cuda.global_avg_pool(xg); yg.backward(g.to(DEV)); _cmp(yg, y); _cmp(xg.grad, x.grad)
    a, b = torch.randn(4, 16, 5, 5), torch.randn(4, 16, 5, 5)
    _cmp(cuda.add_relu(a.to(DEV), b.to(DEV)), F.relu(a + b))
    # dropout: keep-rate, scaling, mask reuse in backward, channel mode
    for per_channel, shape in ((False, (128, 128)), (True, (64, 64, 12, 12)), (False, (64, 64, 12, 12))):
        t = torch.ones(*shape, device=DEV, requires_grad=True)
        d = ops.Dropout2d(0.25) if per_channel else ops.Dropout(0.25)
        d.train(); y = d(t); y.sum().backward()
        keep = (y != 0).float().mean().item()
        assert abs(keep - 0.75) < 0.03, keep
        assert torch.allclose(y[y != 0], torch.tensor(1 / 0.75, device=DEV))
        assert torch.equal(t.grad, y)  # same mask in backward
        if per_channel:
            planes = (y != 0).float().flatten(2)
            assert ((planes.min(2).values == planes.max(2).values).all())  # whole (n, c) planes
        d.eval(); assert torch.equal(d(t), t)


def test_fused_optimizers_match_torch():
    torch.manual_seed(0)
    shapes = [(128, 9216), (128,), (64, 32, 3, 3), (10,)]
    for ours_cls, ref_cls, kw in ((ops.Adadelta, torch.optim.Adadelta, dict(lr=0.5)),
                                  (ops.SGD, torch.optim.SGD, dict(lr=0.1, momentum=0.9, weight_decay=1e-3)),
                                  (ops.Adam, torch.optim.Adam, dict(lr=1e-2))):
        ps = [torch.randn(*s) for s in shapes]
        a = [p.clone().to(DEV).requires_grad_(True) for p in ps]
        b = [p.clone().requires_grad_(True) for p in ps]
        oa, ob = ours_cls(a, **kw), ref_cls(b, **kw)
        sched = ops.StepLR(oa, 1, 0.7)
        for step in range(4):
            for pa, pb in zip(a, b):
                g = torch.randn_like(pb)
                pb.grad = g.clone(); pa.grad = g.to(DEV)
            oa.step(); ob.step()
            if step == 1:
                sched.step()
                for grp in ob.param_groups: grp["lr"] *= 0.7
            for pa, pb in zip(a, b):
                _cmp(pa, pb, rtol=1e-4, atol=1e-5, what=f"{ours_cls.__name__} step {step}")


def test_convnet_forward_backward_matches_cpu():
    """Whole reference model (dropout disabled so both sides see the same graph)."""
    torch.manual_seed(0)
    cpu = dcp.models.ConvNet(); cpu.dropout1.p = 0.0; cpu.dropout2.p = 0.0
    gpu = dcp.models.ConvNet(); gpu.dropout1.p = 0.0; gpu.dropout2.p = 0.0
    gpu.load_state_dict(cpu.state_dict()); gpu.to(DEV)
    x = torch.randn(128, 1, 28, 28); t = torch.randint(0, 10, (128,))
    out = cpu(x); loss = F.nll_loss(out, t); loss.backward()
    og = gpu(x.to(DEV)); lg = ops.nll_loss(og, t.to(DEV)); lg.backward()
    _cmp_norm(og, out, 5e-3, "log-probs"); _cmp(lg, loss, rtol=1e-3, atol=1e-3, what="loss")
    for (n, p), q in zip(gpu.named_parameters(), cpu.parameters()):
        if n == "fc1.bias":   # exactly zero in exact arithmetic (BatchNorm follows fc1)
            assert p.grad.abs().max().item() < 1e-5
        else:
            _cmp_norm(p.grad, q.grad, 6e-2, n)
    assert list(gpu.state_dict().keys()) == list(cpu.state_dict().keys())


@pytest.mark.parametrize("M,C,dtype", [(8192, 3072, torch.bfloat16), (4100, 72, torch.float32), (3000, 768, torch.bfloat16), (100, 10, torch.float32)])
def test_colsum_bias_gradient(M, C, dtype):
    """dcp_colsum (bias gradients of Linear / conv): 16-byte kernel for aligned shapes, scalar kernel otherwise, vs an fp64 sum."""
    from distributed_compute_pytorch_b200.ops import cuda
    L = cuda.lib()
    torch.manual_seed(M + C)
    x = torch.randn(M, C, device=DEV).to(dtype)
    out = torch.full((C,), float("nan"), device=DEV)
    cuda._chk(L.dcp_colsum(x.data_ptr(), out.data_ptr(), M, C, cuda._dt(x), cuda._stream()))
    ref = x.double().sum(0)
    assert ((out.double() - ref).abs().max() / ref.abs().max()).item() < 1e-4
