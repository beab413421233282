import sys, torch
sys.path.insert(0, '/root/repo')
import distributed_compute_pytorch_b200 as dcp
from distributed_compute_pytorch_b200 import ops
from tests.helpers import free_port, init
def worker(rank, world, port):
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    dev = torch.device("cuda", rank)
    g = torch.Generator().manual_seed(3)
    for shape, relu in (((8 * world, 64, 6, 6), True), ((8 * world, 64, 6, 6), False), ((16 * world, 128), False), ((16 * world, 128), True)):
        X = (torch.randn(*shape, generator=g) * 2 + 0.5); G = torch.randn(*shape, generator=g); C = shape[1]
        ref = (ops.BatchNorm2d if len(shape) == 4 else ops.BatchNorm1d)(C, fuse_relu=relu).to(dev)
        sbn = ops.SyncBatchNorm(C, fuse_relu=relu).to(dev); sbn.load_state_dict(ref.state_dict())
        xr = X.to(dev)
        if len(shape) == 4: xr = xr.contiguous(memory_format=torch.channels_last)
        xr.requires_grad_(True)
        yr = ref(xr); yr.backward(G.to(dev))
        n = shape[0] // world
        xl = X[rank * n:(rank + 1) * n].to(dev)
        if len(shape) == 4: xl = xl.contiguous(memory_format=torch.channels_last)
        xl.requires_grad_(True)
        yl = sbn(xl); yl.backward(G[rank * n:(rank + 1) * n].to(dev))
        torch.cuda.synchronize()
        print(rank, shape, relu, 'y', (yl - yr[rank*n:(rank+1)*n]).abs().max().item(), 'dx', (xl.grad - xr.grad[rank*n:(rank+1)*n]).abs().max().item(), 'dx ref max', xr.grad.abs().max().item(), flush=True)
    dist.destroy_process_group()
if __name__ == '__main__':
    dcp.spawn(worker, args=(2, free_port()), nprocs=2, join=True)
