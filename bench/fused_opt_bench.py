"""Isolated timing of dcp_fused_rs_opt_ag (AdamW) on a 32 M-element bucket; world size = number of spawned ranks."""
import ctypes as C, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init

def worker(rank, world, port):
    os.environ.setdefault("DCP_HEAP_MB", "2048")
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    be = dist.distributed._world().cuda
    from distributed_compute_pytorch_b200.ops import cuda as K
    from distributed_compute_pytorch_b200.parallel.p2p_backend import CH_OPT
    L = K.lib(); cl = be.lib
    cl.dcp_fused_shard_elems.restype = C.c_size_t; cl.dcp_fused_shard_elems.argtypes = [C.c_size_t, C.c_int, C.c_int]
    cl.dcp_fused_rs_opt_ag.restype = C.c_int
    cl.dcp_fused_rs_opt_ag.argtypes = [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p]
    n = 32 * 1024 * 1024
    g = be.empty_symmetric(n, torch.float32); p = be.empty_symmetric(n, torch.float32)
    g.normal_(); p.normal_()
    shard = int(cl.dcp_fused_shard_elems(n, world, 0))
    s1 = torch.zeros(shard, device=g.device); s2 = torch.zeros(shard, device=g.device)
    hp = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.01, 1.0, 1.0, 1.0], device=g.device)
    st = torch.cuda.current_stream().cuda_stream
    def run(ctas):
        rc = cl.dcp_fused_rs_opt_ag(be.handle, g.data_ptr(), p.data_ptr(), s1.data_ptr(), s2.data_ptr(), n, hp.data_ptr(), 2, 0, 1, CH_OPT, ctas, st)
        assert rc == 0
    for ctas in (32, 148, 296):
        for _ in range(3): run(ctas)
        dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): run(ctas)
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) * 100
        # per rank: read W grads (local + peers) of its shard, p, s1, s2; write s1, s2 and W copies of p
        hbm = n / world * 4 * (1 + 1 + 2 + 2 + 1)      # local gradient, p read, states r/w, own p write
        link = n / world * 4 * (world - 1) * 2 if world > 1 else 0   # peer gradient pulls + parameter pushes (per direction roughly half each)
        if rank == 0:
            print(f"world {world} ctas {ctas}: {us:.1f} us  HBM {hbm/1e6:.0f} MB -> {hbm/us/1e6:.2f} TB/s  NVLink {link/1e6:.0f} MB -> {link/us/1e3:.0f} GB/s", flush=True)
    dist.destroy_process_group()

if __name__ == "__main__":
    w = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    dcp.spawn(worker, args=(w, free_port()), nprocs=w, join=True)
