"""Fused ConvNet training step (flat-shift tcgen05 convs + fused kernels + CUDA graph) vs the layer-by-layer
autograd path and vs a pure-PyTorch CPU reference of the same model."""
import copy

import pytest
import torch
import torch.nn.functional as F

import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / (b.norm() + 1e-30)).item()


def _make(p1=0.0, p2=0.0):
    torch.manual_seed(0)
    m = dcp.models.ConvNet()
    m.dropout1.p, m.dropout2.p = p1, p2
    return m


@pytest.mark.parametrize("use_graph", [False, True])
def test_gradients_and_update_match_cpu_reference(use_graph):
    N = 128
    cpu = _make()
    gpu = copy.deepcopy(cpu).to(DEV)
    x = torch.randn(N, 1, 28, 28); y = torch.randint(0, 10, (N,))
    # CPU reference: plain torch ops, fp32
    cpu.train()
    ref_opt = torch.optim.Adadelta(cpu.parameters(), lr=0.5)
    loss_ref = F.nll_loss(cpu(x), y); loss_ref.backward()
    ref_grads = {n: p.grad.clone() for n, p in cpu.named_parameters()}
    ref_opt.step()
    opt = ops.Adadelta(gpu.parameters(), lr=0.5)
    gpu.train()
    tr = FusedConvNetTrainer(gpu, opt, N, use_graph=False)
    before = {n: p.detach().clone() for n, p in gpu.named_parameters()}
    if use_graph:
        # capture on a copy of the state, then restore so exactly one update is compared
        state = copy.deepcopy(gpu.state_dict())
        tr._capture()
        gpu.load_state_dict(state)
        for st in opt.state.values():
            for v in st.values():
                if torch.is_tensor(v): v.zero_()
        opt._dcp_cache[0]["hp"][6] = 0.0
    loss = tr.step(x.to(DEV), y.to(DEV))
    torch.cuda.synchronize()
    assert abs(loss.item() - loss_ref.item()) < 2e-3 * max(1.0, abs(loss_ref.item()))
    # tolerance: tf32 multiplies flip a few ReLU masks whose pre-activation is ~0 (a few % in Frobenius norm,
    # the same class of difference cuDNN-tf32 shows against an fp32 reference).  fc1.bias is excluded from the
    # relative check: BatchNorm makes its true gradient exactly zero, so only its absolute size is meaningful.
    for n, p in gpu.named_parameters():
        if n == "fc1.bias":
            assert p.grad.abs().max().item() < 1e-5
            continue
        assert _relerr(p.grad, ref_grads[n]) < 6e-2, (n, _relerr(p.grad, ref_grads[n]))
    # the update: Adadelta's rule (torch/optim/adadelta.py) applied in fp64 to the gradients the kernels produced must give the
    # parameters and the state the fused optimizer kernels left behind (first step: both accumulators start at zero)
    lr, rho, eps = 0.5, 0.9, 1e-6
    for n, p in gpu.named_parameters():
        g = p.grad.double()
        sq = (1 - rho) * g * g
        delta = torch.sqrt(torch.full_like(g, eps)) / torch.sqrt(sq + eps) * g
        exp_p = before[n].double() - lr * delta
        assert torch.allclose(p.detach().double(), exp_p, rtol=1e-5, atol=1e-6), (n, (p.detach().double() - exp_p).abs().max().item())
        st = opt.state[p]
        assert torch.allclose(st["square_avg"].double(), sq, rtol=1e-5, atol=1e-12), n
        assert torch.allclose(st["acc_delta"].double(), (1 - rho) * delta * delta, rtol=1e-4, atol=1e-12), n
        assert torch.isfinite(p).all()
    assert _relerr(gpu.batchnorm.running_mean, cpu.batchnorm.running_mean) < 5e-3
    assert _relerr(gpu.batchnorm.running_var, cpu.batchnorm.running_var) < 5e-3
    assert gpu.batchnorm.num_batches_tracked.item() == 1


def test_training_reduces_loss_with_dropout_and_graph():
    N = 128
    torch.manual_seed(1)
    net = dcp.models.ConvNet().to(DEV).train()
    opt = ops.Adadelta(net.parameters(), lr=1.0)
    tr = FusedConvNetTrainer(net, opt, N, use_graph=True)
    assert tr.launches_per_step > 0
    ds = dcp.utils.SyntheticMNIST(2048, seed=3)
    xs, ys = ds.images.to(DEV), ds.labels.to(DEV)
    losses = []
    for ep in range(6):
        for i in range(0, 2048, N):
            losses.append(tr.step(xs[i:i + N], ys[i:i + N]).item())
    assert all(map(lambda v: v == v, losses))  # no NaN
    assert sum(losses[-8:]) / 8 < 0.7 * sum(losses[:8]) / 8, (losses[:4], losses[-4:])
    # dropout masks change from replay to replay (device-side step counter)
    tr.x.copy_(xs[:N]); tr.y.copy_(ys[:N])
    tr.graph.replay(); a = tr.pooled.clone(); tr.graph.replay(); b = tr.pooled.clone()
    assert not torch.equal(a == 0, b == 0)


def test_fused_head_kernel_matches_kernel_per_op_head():
    """dcp_convnet_head (one CTA: BN+ReLU+dropout+fc2+CE forward and backward) vs the separate kernels, with dropout
    ON: both derive their masks from the same Philox (seed, stream, step, index), so results must agree closely."""
    N = 128
    outs = []
    for fused in (True, False):
        torch.manual_seed(3)
        net = dcp.models.ConvNet().to(DEV).train()
        opt = ops.Adadelta(net.parameters(), lr=0.0)
        tr = FusedConvNetTrainer(net, opt, N, use_graph=False, seed=1234)
        tr.fused_head = fused
        from distributed_compute_pytorch_b200.ops import cuda as K
        K.graph_step.zero_()
        g = torch.Generator().manual_seed(9)
        x = torch.randn(N, 1, 28, 28, generator=g).to(DEV); y = torch.randint(0, 10, (N,), generator=g).to(DEV)
        loss = tr.step(x, y).clone()
        torch.cuda.synchronize()
        outs.append((loss, tr.dh1.clone(), {n: p.grad.clone() for n, p in net.named_parameters()},
                     net.batchnorm.running_mean.clone(), net.batchnorm.running_var.clone()))
    (l0, d0, g0, rm0, rv0), (l1, d1, g1, rm1, rv1) = outs
    assert abs(l0.item() - l1.item()) < 1e-4
    assert _relerr(d0, d1) < 1e-3
    assert _relerr(rm0, rm1) < 1e-5 and _relerr(rv0, rv1) < 1e-4
    for n in g0:
        if n != "fc1.bias":
            assert _relerr(g0[n], g1[n]) < 2e-3, (n, _relerr(g0[n], g1[n]))
    # fc1.bias feeds a BatchNorm: its exact gradient is 0, both paths leave rounding noise only
    assert g0["fc1.bias"].abs().max().item() < 1e-4 and g1["fc1.bias"].abs().max().item() < 1e-4


def test_split_optimizer_and_steplr_across_graph_replays():
    """The captured step updates bucket 0 on the side stream and bucket 1 at the end (two launches over slices of the
    optimizer's pointer table) and reads lr from device memory.  Against the eager, single-launch optimizer path, with
    StepLR changing the learning rate between steps, the parameters must agree to rounding."""
    N = 128
    g = torch.Generator().manual_seed(5)
    xs = torch.randn(4, N, 1, 28, 28, generator=g).to(DEV); ys = torch.randint(0, 10, (4, N), generator=g).to(DEV)
    results = []
    for graph in (False, True):
        net = _make().to(DEV).train()
        opt = ops.Adadelta(net.parameters(), lr=1.0)
        sched = ops.StepLR(opt, step_size=1, gamma=0.5)
        tr = FusedConvNetTrainer(net, opt, N, use_graph=False)
        tr.split_opt = graph
        if graph:
            state = copy.deepcopy(net.state_dict())
            tr._capture()
            net.load_state_dict(state)
            for st in opt.state.values():
                for v in st.values():
                    if torch.is_tensor(v): v.zero_()
            opt._dcp_cache[0]["hp"][6] = 0.0
            assert tr._opt_rows() is not None
        for i in range(4):
            tr.step(xs[i], ys[i])
            sched.step()
        torch.cuda.synchronize()
        results.append({n: p.detach().clone() for n, p in net.named_parameters()})
    for n in results[0]:
        assert _relerr(results[1][n], results[0][n]) < 1e-5, (n, _relerr(results[1][n], results[0][n]))


@pytest.mark.parametrize("N,training", [(32, 1), (128, 1), (64, 0)])
def test_conv2_with_fused_pool_epilogue_matches_conv_then_pool(N, training):
    """dcp_conv_flat_pool_tc (conv2 + bias + ReLU + 2x2 max-pool + channel dropout in the tcgen05 epilogue) against the two-kernel
    path (dcp_conv_flat_tc then dcp_convnet_pool_fwd): same accumulators, same Philox stream -> identical pooled values and
    argmax bytes; and both against a plain PyTorch fp32 conv / relu / max_pool2d of the same input (tf32 tolerance)."""
    import ctypes as C
    from distributed_compute_pytorch_b200.ops import cuda as K
    L = K.lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(N)
    x = torch.randn(N, 26, 26, 32, device=dev, generator=g)                      # flat pixels [N*676, 32] (NHWC)
    w = torch.randn(64, 3, 3, 32, device=dev, generator=g) * 0.1                 # [Cout][taps*Cin], tap-major
    b = torch.randn(64, device=dev, generator=g) * 0.1
    off = (C.c_int * 9)(*[r * 26 + s for r in range(3) for s in range(3)])
    step = torch.full((), 3, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    a2 = torch.empty(N * 676, 64, device=dev)
    p_ref = torch.full((N, 9216), float("nan"), device=dev); a_ref = torch.full((N * 9216,), 255, dtype=torch.uint8, device=dev)
    p_new = torch.full((N, 9216), float("nan"), device=dev); a_new = torch.full((N * 9216,), 255, dtype=torch.uint8, device=dev)
    assert L.dcp_conv_flat_tc(x.data_ptr(), w.data_ptr(), a2.data_ptr(), b.data_ptr(), None, N * 676, 32, 64, 9, off, 26, 676, 24, 24, 0, 148, st) == 0
    assert L.dcp_convnet_pool_fwd(a2.data_ptr(), p_ref.data_ptr(), a_ref.data_ptr(), N, 0.25, 1234, 1, step.data_ptr(), training, st) == 0
    rc = L.dcp_conv_flat_pool_tc(x.data_ptr(), w.data_ptr(), p_new.data_ptr(), a_new.data_ptr(), b.data_ptr(), N * 676, 32, 64, 9, off, 26, 676, 24, 24,
                                 0.25, 1234, 1, step.data_ptr(), training, 148, st)
    assert rc == 0, L.dcp_last_error()
    torch.cuda.synchronize()
    assert torch.equal(p_new, p_ref), (p_new - p_ref).abs().max().item()
    assert torch.equal(a_new, a_ref)
    if not training:
        ref = torch.nn.functional.max_pool2d(torch.relu(torch.nn.functional.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), b)), 2)
        assert torch.allclose(p_new.view(N, 64, 12, 12), ref, rtol=2e-2, atol=2e-2)
