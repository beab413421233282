"""Small single-GPU programs whose kernels are worth one `ncu --set full` capture each (profiles/README.md lists the commands).

    python bench/ncu_targets.py coll     # collective bodies in a 4-rank emulated world (k_coll_emu<...>: all_reduce two-shot / one-shot,
                                         # all_gather, reduce_scatter, broadcast, fused reduce_scatter -> AdamW -> all_gather), 16 MiB fp32
    python bench/ncu_targets.py ops      # BatchNorm forward / backward, cross-entropy forward / backward, multi-tensor Adadelta + AdamW

The emulated world runs every rank of a collective inside one cooperative launch on this GPU, so the protocol and the reduction
loops are the production ones while "peer" traffic stays on the device: read the captures for instruction mix, stalls and local
memory behaviour, not for NVLink bandwidth (bench/coll_sweep.py measures that on real peers).
"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def coll():
    from distributed_compute_pytorch_b200.distributed import ReduceOp as R
    from distributed_compute_pytorch_b200.parallel.emulated import EmulatedWorld
    from distributed_compute_pytorch_b200.parallel.p2p_backend import CH_OPT
    world, n = 4, 4 << 20                      # 16 MiB fp32 per rank
    dev = torch.device("cuda", 0)
    with EmulatedWorld(world, heap_mb=256) as w:
        ts = [be.empty_symmetric(n, torch.float32) for be in w.ranks]
        outs = [be.empty_symmetric(n, torch.float32) for be in w.ranks]
        small = [be.empty_symmetric(2048, torch.float32) for be in w.ranks]
        for t in ts + outs + small:
            t.normal_()
        for _ in range(2):
            for r, be in enumerate(w.ranks):
                be.all_reduce(ts[r], R.SUM, algo=2, max_ctas=8)
            w.flush()
            for r, be in enumerate(w.ranks):
                be.all_reduce(small[r], R.SUM, algo=1, max_ctas=8)
            w.flush()
            for r, be in enumerate(w.ranks):
                be.all_gather_into_tensor(outs[r], ts[r][: n // world])
            w.flush()
            for r, be in enumerate(w.ranks):
                be.reduce_scatter_tensor(outs[r][: n // world], ts[r], R.SUM)
            w.flush()
            for r, be in enumerate(w.ranks):
                be.broadcast(ts[r], 1)
            w.flush()
        lib = w.lib
        lib.dcp_fused_shard_elems.restype = C.c_size_t; lib.dcp_fused_shard_elems.argtypes = [C.c_size_t, C.c_int, C.c_int]
        lib.dcp_fused_rs_opt_ag.restype = C.c_int
        lib.dcp_fused_rs_opt_ag.argtypes = [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p] + [C.c_int] * 5 + [C.c_void_p]
        shard = lib.dcp_fused_shard_elems(n, world, 0)
        s1 = [torch.zeros(shard, device=dev) for _ in range(world)]; s2 = [torch.zeros(shard, device=dev) for _ in range(world)]
        hp = torch.tensor([1e-3, 0.9, 0.999, 1e-8, 0.01, 1.0, 1.0, 1.0], device=dev)
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            for r, be in enumerate(w.ranks):
                assert lib.dcp_fused_rs_opt_ag(be.handle, ts[r].data_ptr(), outs[r].data_ptr(), s1[r].data_ptr(), s2[r].data_ptr(), n, hp.data_ptr(),
                                               2, 0, 0, CH_OPT, 8, st) == 0
            w.flush()
        torch.cuda.synchronize()
        w.check_health()
    print("ok")


def ops_():
    import torch.nn.functional as F
    from distributed_compute_pytorch_b200 import ops
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    # BatchNorm (+ReLU) on a ResNet-50 sized activation, bf16 channels-last
    bn = ops.BatchNorm2d(256).to(dev)
    x = torch.randn(64, 256, 56, 56, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    for _ in range(2):
        y = bn(x)
        y.float().square().mean().backward()
    # cross-entropy on GPT-2 sized logits
    logits = torch.randn(8192, 50304, device=dev, dtype=torch.bfloat16, requires_grad=True)
    tgt = torch.randint(0, 50257, (8192,), device=dev)
    for _ in range(2):
        ops.cross_entropy(logits, tgt).backward()
    # multi-tensor optimizers over ~25M parameters in 60 tensors
    ps = [torch.randn(640, 640, device=dev, requires_grad=True) for _ in range(60)]
    for opt in (ops.Adadelta(ps, lr=1.0), ops.Adam(ps, lr=1e-3, weight_decay=0.1, decoupled=True)):
        for _ in range(2):
            for p in ps:
                p.grad = torch.ones_like(p)
            opt.step()
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    {"coll": coll, "ops": ops_}[sys.argv[1] if len(sys.argv) > 1 else "coll"]()
