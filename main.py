"""Reference-compatible DDP MNIST example on the B200-native framework.

Same flags, prints, loops and checkpoint as /root/reference/main.py (ref:main.py:55-156), running on
``distributed_compute_pytorch_b200`` instead of torch.distributed / torch DDP.  Deliberate deviations
(SURVEY.md section 9): CPU runs use ``torch.device('cpu')`` instead of crashing on ``.to(rank)``; the
model is DDP-wrapped whenever world_size > 1; only rank 0 writes ``mnist.pt`` (followed by a barrier);
``--no-cuda`` really selects the CPU; data is synthetic MNIST-shaped unless ``--data mnist`` finds
the files locally (there is no network to download them, ref:main.py:107).
"""
import argparse
import os
import time

import torch
from torch.utils.data import DataLoader

import distributed_compute_pytorch_b200 as dist
from distributed_compute_pytorch_b200 import DistributedDataParallel, DistributedSampler, ops
from distributed_compute_pytorch_b200.models import ConvNet
from distributed_compute_pytorch_b200.utils import SyntheticMNIST


def setup(rank, world_size, backend=None):
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")   # ref:main.py:48 uses "localhost"
    os.environ.setdefault("MASTER_PORT", "12355")       # ref:main.py:49
    dist.init_process_group(backend, rank=rank, world_size=world_size)


def cleanup():
    dist.destroy_process_group()


class _LossPrinter:
    """The reference's every-10th-batch line (ref:main.py:64-68: all_reduce(SUM) of the batch loss, printed on rank 0) without
    stalling the GPU: the reduced value travels to pinned host memory behind an event and is printed once it has arrived - at
    the next print step or at the end of the pass - in the same order and with the same text."""

    def __init__(self, rank, device):
        self.rank, self.pending = rank, []
        self.cuda = torch.device(device).type == "cuda" if not isinstance(device, int) else True

    def submit(self, loss, epoch, b, total):
        red = loss.detach().clone()
        dist.all_reduce(red, op=dist.ReduceOp.SUM)
        if self.rank != 0:
            return
        if not self.cuda:
            self._emit(epoch, b, total, red.item()); return
        host = torch.empty((), dtype=red.dtype).pin_memory()
        host.copy_(red, non_blocking=True)
        ev = torch.cuda.Event(); ev.record()
        self.pending.append((ev, host, epoch, b, total))
        self.drain(block=False)

    def drain(self, block=True):
        while self.pending and (block or self.pending[0][0].query()):
            ev, host, epoch, b, total = self.pending.pop(0)
            ev.synchronize()
            self._emit(epoch, b, total, host.item())

    @staticmethod
    def _emit(epoch, b, total, value):
        print(f"epoch: {epoch} [{b}/{total} \
          ({100. * b/total}%)]\t Loss:{value}")


def _make_engine(opt, model, optimizer, device):
    """--engine: 'eager' = layer-by-layer autograd (the reference's loop body), 'graph' = the same body captured once in a CUDA
    graph (any model), 'fused' = the hand-scheduled flagship step (parallel.FusedConvNetTrainer: ~19 kernels, one graph,
    gradient exchange fused into the kernels).  'auto' picks the fastest one the run is eligible for."""
    dev = torch.device(device) if not isinstance(device, int) else torch.device("cuda", device)
    want = opt.engine
    if dev.type != "cuda" or want == "eager":
        return None
    net = model.module if hasattr(model, "module") else model
    fused_ok = isinstance(net, ConvNet) and opt.batch_size % 32 == 0 and 32 <= opt.batch_size <= 128
    if want in ("auto", "fused") and fused_ok:
        from distributed_compute_pytorch_b200.parallel.convnet_trainer import FusedConvNetTrainer
        return FusedConvNetTrainer(model, optimizer, opt.batch_size)
    if want == "fused":
        raise SystemExit("--engine fused needs the ConvNet model and a batch size that is a multiple of 32 in [32, 128]")
    from distributed_compute_pytorch_b200.parallel.engine import GraphedTrainStep
    x = torch.zeros(opt.batch_size, 1, 28, 28, device=dev); y = torch.zeros(opt.batch_size, dtype=torch.int64, device=dev)
    return GraphedTrainStep(model, optimizer, lambda o, t: ops.nll_loss(o, t), x, y, device=dev)


def train(opt, model, rank, train_loader, optimizer, epoch, device=None, engine=None):
    device = rank if device is None else device
    model.train()
    printer = _LossPrinter(rank, device)
    for b, (img, label) in enumerate(train_loader):
        if engine is not None and len(img) == opt.batch_size:
            loss = engine(img, label)          # pinned host batch -> staged H2D copy -> one graph replay
        else:                                  # eager path: also takes the (smaller) last batch of a pass
            img, label = img.to(device), label.to(device)
            optimizer.zero_grad()
            out = model(img)
            loss = ops.nll_loss(out, label)
            loss.backward()
            optimizer.step()
        if b % 10 == 0:
            printer.submit(loss, epoch, b, len(train_loader))
        if opt.max_steps and b + 1 >= opt.max_steps:
            break
    printer.drain()


def test(opt, model, rank, test_loader, optimizer, epoch, device=None):
    device = rank if device is None else device
    model.eval()
    test_loss_tensor = torch.zeros((), device=device)
    correct_tensor = torch.zeros((), dtype=torch.int64, device=device)
    datalen = 0
    with torch.no_grad():
        for b, (img, label) in enumerate(test_loader):
            img, label = img.to(device), label.to(device)
            out = model(img)
            # device-side accumulation: no per-batch host sync (the reference does two, ref:main.py:81-85)
            test_loss_tensor += ops.nll_loss(out, label, reduction='sum')
            correct_tensor += ops.correct_count(out, label)
            datalen += len(img)
            if opt.max_steps and b + 1 >= opt.max_steps:
                break
    dist.all_reduce(test_loss_tensor, op=dist.ReduceOp.SUM)
    dist.all_reduce(correct_tensor, op=dist.ReduceOp.SUM)
    if rank == 0:
        n = len(test_loader.dataset)
        print('\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n'.format(
            test_loss_tensor.item(), correct_tensor.item(), n, 100. * correct_tensor.item() / n))
    return test_loss_tensor, correct_tensor


def make_dataset(opt, train=True):
    if opt.data == "mnist":
        from torchvision import datasets, transforms
        return datasets.MNIST('./data/', train=train, download=False, transform=transforms.Compose(
            [transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))]))
    return SyntheticMNIST(opt.data_size if train else max(opt.data_size // 6, 1), seed=1234 if train else 4321)


def proc(rank, world_size, opt, use_cuda):
    print(f"running DDP example on rank {rank}")
    setup(rank, world_size, opt.backend)

    torch.manual_seed(0)
    device = torch.device('cuda', rank % max(torch.cuda.device_count(), 1)) if use_cuda else torch.device('cpu')
    if use_cuda:
        torch.cuda.set_device(device)

    train_dataset = make_dataset(opt, True)
    train_sampler = DistributedSampler(train_dataset, num_replicas=world_size, rank=rank)
    train_loader = DataLoader(train_dataset, batch_size=opt.batch_size, shuffle=False,
                              num_workers=0, pin_memory=use_cuda, sampler=train_sampler)
    test_dataset = make_dataset(opt, False)
    test_sampler = DistributedSampler(test_dataset, num_replicas=world_size, rank=rank)
    test_loader = DataLoader(test_dataset, batch_size=opt.batch_size, shuffle=False, sampler=test_sampler)  # noqa: F841 (unused in the reference too)

    model = ConvNet().to(device)
    if world_size > 1:
        model = DistributedDataParallel(model, device_ids=[device.index] if use_cuda else None)

    optimizer = ops.Adadelta(model.parameters(), lr=opt.lr)
    scheduler = ops.StepLR(optimizer, step_size=1, gamma=opt.gamma)

    from distributed_compute_pytorch_b200.utils import Tracer, load_model, load_training_state, save_model, save_training_state
    tracer = Tracer(rank=rank)                 # DCP_TRACE=<prefix> writes <prefix>_rank<r>.json (chrome://tracing) at the end
    if opt.no_dropout:
        net = model.module if hasattr(model, "module") else model
        net.dropout1.p = net.dropout2.p = 0.0
    start_epoch = 0
    if opt.resume and os.path.exists(opt.checkpoint):      # not in the reference (save-only): model file + optimizer / schedule / epoch sidecar
        load_model(opt.checkpoint, model)
        if os.path.exists(opt.checkpoint + ".state"):
            start_epoch = int(load_training_state(opt.checkpoint + ".state", optimizer, scheduler).get("epoch") or 0)
        if rank == 0:
            print(f"resumed from {opt.checkpoint} at epoch {start_epoch}")
    engine = _make_engine(opt, model, optimizer, device)      # built after --resume: it captures the optimizer's state pointers
    for epoch in range(start_epoch, opt.epochs):
        t0 = time.time()
        with tracer.range(f"train epoch {epoch}"):
            train(opt, model, rank, train_loader, optimizer, epoch, device, engine)
        with tracer.range(f"eval epoch {epoch}"):
            test(opt, model, rank, train_loader, optimizer, epoch, device)   # the reference evaluates on the train loader (ref:main.py:130)
        scheduler.step()
        print(f"time to complete this epoch: {time.time() - t0} seconds")
    if tracer.enabled:
        tracer.export_chrome(f"{os.environ['DCP_TRACE']}_rank{rank}.json")
    # rank 0 only, temp file + rename, then a barrier (every rank writes the same file in the reference, ref:main.py:133).  The
    # optimizer / schedule / epoch sidecar goes first, so a model file never sits next to an older sidecar.
    save_training_state(opt.checkpoint + ".state", optimizer, scheduler, epoch=opt.epochs)
    save_model(opt.checkpoint, model)
    cleanup()


def _str2bool(v):
    return str(v).lower() in ("1", "true", "yes", "y", "")


def build_parser():
    parser = argparse.ArgumentParser()
    parser.add_argument('--batch_size', type=int, default=128, help="batch size of train and test")
    parser.add_argument('--lr', type=float, default=0.001, help="LR of optimizer")
    parser.add_argument('--epochs', type=int, default=20, help="#of epochs")
    parser.add_argument('--no-cuda', nargs='?', const=True, default=False, type=_str2bool, help="disables CUDA if True")
    parser.add_argument('--gamma', default=0.7, type=float, help="gamma value for lr update")
    parser.add_argument('--gpus', default=4, type=int, help="# of GPUs")
    # additions (not in the reference)
    parser.add_argument('--data', default="synthetic", choices=["synthetic", "mnist"])
    parser.add_argument('--data-size', type=int, default=60000, help="synthetic train-set size")
    parser.add_argument('--backend', default=None, help="gloo|nccl|b200|cpu (default: GPU backend iff CUDA)")
    parser.add_argument('--max-steps', type=int, default=0, help="stop each epoch pass after this many batches (0 = all)")
    parser.add_argument('--checkpoint', default="mnist.pt")
    parser.add_argument('--engine', default="auto", choices=["auto", "eager", "graph", "fused"],
                        help="training-step engine on CUDA: eager autograd, the same step in one CUDA graph, or the fused flagship step")
    parser.add_argument('--no-dropout', action='store_true', help="p = 0 for both dropout layers (deterministic runs / engine comparisons)")
    parser.add_argument('--resume', action='store_true', help="continue from --checkpoint (+ .state sidecar) and write the sidecar at the end")
    return parser


def main(argv=None):
    opt = build_parser().parse_args(argv)
    use_cuda = not opt.no_cuda and torch.cuda.is_available()
    world_size = min(opt.gpus, torch.cuda.device_count()) if use_cuda else 2
    dist.spawn(proc, args=(world_size, opt, use_cuda), nprocs=world_size, join=True)


if __name__ == '__main__':
    main()
