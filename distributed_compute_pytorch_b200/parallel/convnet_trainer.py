"""Fused training step for the flagship model (the reference's ConvNet, ref:main.py:20-45, trained by the
loop at ref:main.py:55-63): forward, loss, backward, bucketed gradient all-reduce and Adadelta update
issued as ~30 hand-written kernels with no autograd bookkeeping in between, captured in ONE CUDA graph.

Design notes
* Activations of the convolutional trunk live in the flat 26x26 frame of conv1's output
  (csrc/ops/conv_flat_tc.cu): conv2 forward, its data-gradient and its weight-gradient are shifted GEMMs on
  tcgen05 with the filter resident in shared memory and each pixel tile loaded once.
* Gradients are written straight into the DDP bucket views (no copy-in); bucket 0 (fc2, batchnorm, fc1 -
  4.7 MB) is all-reduced on the side stream while the convolution backward is still running, bucket 1
  (conv2, conv1 - 75 KB) right after - the same order and sizes as the library reducer uses for this model.
* Every hyper-parameter the kernels need (lr, step, dropout stream) is read from device memory, so StepLR and
  fresh dropout masks work across graph replays.
Semantics are those of ``loss = nll_loss(model(x), y); loss.backward(); optimizer.step()`` on the same
weights (tests/test_convnet_trainer_gpu.py checks gradients against the layer-by-layer autograd path).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from .. import distributed as dist
from ..models.convnet import ConvNet
from ..ops import cuda as K
from ..ops import tc
from ..types import ReduceOp
from .ddp import DistributedDataParallel
from .p2p_backend import CH_DDP, CH_TAIL

READY_ORDER = ["fc2.bias", "fc2.weight", "batchnorm.weight", "batchnorm.bias", "fc1.bias", "fc1.weight",
               "conv2.weight", "conv2.bias", "conv1.weight", "conv1.bias"]


class FusedConvNetTrainer:
    def __init__(self, model, optimizer, batch_size: int, use_graph: bool = True, seed: Optional[int] = None):
        self.ddp = model if isinstance(model, DistributedDataParallel) else None
        self.net: ConvNet = model.module if self.ddp is not None else model
        if not isinstance(self.net, ConvNet):
            raise TypeError("FusedConvNetTrainer drives models.ConvNet (optionally DDP-wrapped)")
        self.opt, self.N = optimizer, int(batch_size)
        if self.N < 32 or self.N % 32:
            raise ValueError(f"FusedConvNetTrainer: batch size {self.N} must be a multiple of 32 (the batch is the K axis of fc1's weight-gradient GEMM); "
                             "use GraphedTrainStep or the plain autograd path for other sizes")
        net, N = self.net, self.N
        dev = net.fc1.weight.device
        self.dev = dev
        K.lib()
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.be = dist._world().cuda if (dist.is_initialized() and self.world > 1) else None
        names = dict(net.named_parameters())
        # ---- gradient storage: DDP bucket views (rebuilt in the known ready order) or one private flat buffer
        if self.ddp is not None:
            pos = {id(q): i for i, q in enumerate(self.ddp._params)}
            order = [pos[id(names[n])] for n in READY_ORDER]
            self.ddp._build_buckets(order, limits=[self.ddp.first_bucket_bytes, self.ddp.bucket_bytes_cap])
            self.ddp._rebuilt = True
            for i, p in enumerate(self.ddp._params):
                p.grad = self.ddp._view_of[i]
            self.buckets = [b.flat for b in self.ddp._buckets]
        else:
            total = sum(-(-names[n].numel() // 4) * 4 for n in READY_ORDER)
            flat = torch.zeros(total, dtype=torch.float32, device=dev)
            off = 0
            for n in READY_ORDER:
                p = names[n]
                v = flat[off: off + p.numel()]
                p.grad = v.view(p.shape) if p.is_contiguous() else v.as_strided(p.shape, p.stride())
                off += -(-p.numel() // 4) * 4
            self.buckets = [flat]
        f32 = dict(dtype=torch.float32, device=dev)
        # ---- BatchNorm buffers in ONE flat block, so the per-forward buffer broadcast DDP does from rank 0
        # (torch semantics, broadcast_buffers=True) is a single small collective on the side stream
        self.bn_flat = None
        bn = net.batchnorm
        if self.be is not None and self.ddp is not None and self.ddp.broadcast_buffers and bn.track_running_stats:
            flat = self.be.empty_symmetric(258, torch.float32)    # symmetric: rank 0's head kernel writes every rank's copy
            flat[:128].copy_(bn.running_mean); flat[128:256].copy_(bn.running_var)
            nbt = flat[256:258].view(torch.int64)
            nbt.copy_(bn.num_batches_tracked.reshape(1))
            bn.running_mean.data, bn.running_var.data, bn.num_batches_tracked.data = flat[:128], flat[128:256], nbt.reshape(())
            self.bn_flat = flat
        P = N * 676
        self.x = torch.zeros(N, 1, 28, 28, **f32)
        self.y = torch.zeros(N, dtype=torch.int64, device=dev)
        self.a1 = torch.empty(P, 32, **f32)
        self.a2 = torch.empty(P, 64, **f32)
        self.pooled = torch.empty(N, 9216, **f32)
        self.arg = torch.empty(N * 9216, dtype=torch.uint8, device=dev)
        self.h1 = torch.empty(N, 128, **f32); self.h2 = torch.empty(N, 128, **f32); self.h3 = torch.empty(N, 128, **f32)
        self.logits = torch.empty(N, 10, **f32)
        self.loss = torch.zeros((), **f32); self.lse = torch.empty(N, **f32); self.one = torch.ones((), **f32)
        self.bn_scratch = torch.empty(4 * 128 + 2, dtype=torch.float64, device=dev)
        self.bn_mean = torch.empty(128, **f32); self.bn_rstd = torch.empty(128, **f32)
        self.dlogits = torch.empty(N, 10, **f32); self.dh3 = torch.empty(N, 128, **f32); self.dh2 = torch.empty(N, 128, **f32)
        self.dh1 = torch.empty(N, 128, **f32); self.dpooled = torch.empty(N, 9216, **f32)
        self.dy2 = torch.empty(P, 64, **f32); self.dz1 = torch.empty(P, 32, **f32)
        self.wt2 = torch.empty(32 * 9 * 64, **f32)
        self.wg_parts = min(148, -(-P // 128))
        self.wg_scratch = torch.empty(self.wg_parts * 64 * 289, **f32)
        self.flat_wgrad = os.environ.get("DCP_FLAT_WGRAD", "1") == "1"
        # filter flip, dropout step counter and optimizer step counter on the side stream at the start of the step.  Default: single GPU
        # only (0.1296 -> 0.1251 ms); with two GPUs, where the side stream is the communication stream, the one measurement taken was
        # slower (0.143 -> 0.151 ms), so the data-parallel step keeps them on the main stream unless DCP_EARLY_SIDE=1
        self.early_side = os.environ.get("DCP_EARLY_SIDE", "1" if self.be is None else "0") == "1"
        self.fc1_overlap = os.environ.get("DCP_FC1_OVERLAP", "1") == "1"     # fc1 weight gradient on the side stream, next to the data gradient
        self.pool_fused = os.environ.get("DCP_POOL_FUSED", "1") == "1"
        self.off_fwd = (C.c_int * 9)(*[r * 26 + s for r in range(3) for s in range(3)])
        self.off_bwd = (C.c_int * 9)(*[-((2 - r) * 26 + (2 - s)) for r in range(3) for s in range(3)])
        self.seed = int(torch.initial_seed() if seed is None else seed) & 0xFFFFFFFFFFFFFFFF
        if K.graph_step is None:
            K.graph_step = torch.zeros((), dtype=torch.int64, device=dev)
        self.training = True
        self.fused_head = True
        self.comm_ctas = 16
        self.split_opt = os.environ.get("DCP_SPLIT_OPT", "1") == "1"
        # data-parallel tail: conv1's 320 gradients are averaged across the ranks INSIDE the conv1 weight-gradient kernel (its last
        # CTA pushes / signals / sums over NVLink peer memory) instead of by a separate all-reduce launch on the side stream
        self.fuse_tail = self.be is not None and os.environ.get("DCP_FUSE_TAIL", "1") == "1"
        self.c1_payload = self.be.empty_symmetric(2 * self.world * 320, torch.float32) if self.fuse_tail else None
        self.side = self.be.comm_stream if self.be is not None else torch.cuda.Stream(device=dev)
        # measurement switches (bench/step_ablation.py): drop one piece of the data-parallel machinery to see what it costs
        # on the critical path.  The step is numerically WRONG with any of them set.
        self.ablate = {a for a in os.environ.get("DCP_ABLATE", "").split(",") if a}
        self.comm_ctas = int(os.environ.get("DCP_COMM_CTAS", self.comm_ctas))
        # fused data-parallel tail (csrc/ops/dp_push.cuh): the producers of the last bucket's gradients (conv2 split-K reduce,
        # conv1 weight-gradient) push their sums into every rank's payload block and signal; ONE kernel at the end of the step
        # waits, averages and applies the optimizer update.  No all-reduce launch, no side stream, no join for that bucket.
        self.dp_tail = (self.be is not None and self.ddp is not None and len(self.buckets) == 2 and self.flat_wgrad
                        and os.environ.get("DCP_DP_TAIL", "1") == "1")
        if self.dp_tail:
            g = {n: q.grad for n, q in names.items()}
            self.dp_tail = self._bucket1_split(g) > 0 and g["conv1.bias"].data_ptr() == g["conv1.weight"].data_ptr() + 4 * 288
        if self.dp_tail:
            self.dp_slot = self.buckets[1].numel()
            self.dp_pay = self.be.empty_symmetric(2 * self.world * self.dp_slot, torch.float32)
            self.dp_pay.zero_()
            self.dp_ws = torch.zeros(16 + 148 * 320, **f32)
            self.c1_ll = self.be.empty_symmetric(2 * self.world * 320 * 2, torch.float32)     # {value, step} words of conv1's exchange
            self.c1_ll.zero_()
            self.dp_blks = (C.c_int * 2)(0, 1)
            self.side2 = torch.cuda.Stream(device=dev, priority=-1)
            # every rank must see the same step number inside a step: the epoch of the exchange is the device step counter
            chk = torch.stack([K.graph_step.double(), -K.graph_step.double()])
            dist.all_reduce(chk, op=ReduceOp.MAX)
            if chk[0].item() != -chk[1].item():
                raise RuntimeError("FusedConvNetTrainer: the ranks disagree on the device step counter (they must have run the same "
                                   "number of fused steps before this trainer was built)")
        if self.be is not None:
            dist.barrier()          # payload / buffer blocks are zeroed everywhere before any rank pushes into them
        self.launches_per_step = 0
        self.graph = None
        self._copy_stream = None
        if use_graph:
            self._capture()

    # ---------------------------------------------------------------------------------------------
    def _issue(self) -> None:
        """Enqueue one full training step on the current stream (and the comm stream for the buckets)."""
        L, s, net, N = K.lib(), K._stream, self.net, self.N
        ck = K._chk
        p = {n: t for n, t in net.named_parameters()}
        g = {n: t.grad for n, t in p.items()}
        step_ptr = K.graph_step.data_ptr()
        tr = int(self.training)
        ab = self.ablate
        self._conv2_tail = None
        main = torch.cuda.current_stream(self.dev)
        rows = self._opt_rows() if self.split_opt else None
        dp_tail = self.dp_tail and rows is not None and "b1" not in ab
        bn = net.batchnorm
        bn_push = self.bn_flat is not None and tr and N <= 128 and self.fused_head and "bn" not in ab and "oldbn" not in ab
        ev_buf = None
        if self.bn_flat is not None and tr and not bn_push and "bn" not in ab:
            # DDP forward semantics: rank 0's buffers win.  Runs on the side stream under conv1 / conv2.
            self.side.wait_stream(main)
            self.be.broadcast(self.bn_flat, 0, stream=self.side, chan=CH_DDP)
            ev_buf = torch.cuda.Event(); ev_buf.record(self.side)
        # Two tiny launches that nothing in the forward pass needs - the data-gradient's copy of conv2's filter (taps flipped, channels
        # swapped; conv2.weight may then be updated while the data-gradient is still running) and the optimizer's step counter - go
        # to the side stream right away instead of sitting between the big kernels of the main stream; main picks them up with an
        # event wait long after they finished.
        ev_early = ev_ctr = None
        early = self.early_side and rows is not None and "opt0main" not in ab and "opt0end" not in ab and ev_buf is None
        # (the dropout step counter stays at the head of the main stream: moved to the side stream as well, conv2 had to wait for an
        # event between conv1 and itself, which breaks their programmatic dependent launch - 0.1251 -> 0.1305 ms)
        ck(L.dcp_counter_advance(step_ptr, s()))
        if early:
            self.side.wait_stream(main)
            with torch.cuda.stream(self.side):
                ck(tc._lib.dcp_flip_weights(p["conv2.weight"].data_ptr(), self.wt2.data_ptr(), 64, 3, 3, 32, 0, s()))
                K._push_hyper(self.opt, self.opt.param_groups[0], rows[0])
                ck(L.dcp_optim_advance(rows[0]["hp"].data_ptr(), s()))
                ev_early = torch.cuda.Event(); ev_early.record(self.side)
        # ---- forward
        ck(L.dcp_convnet_conv1_fwd(self.x.data_ptr(), p["conv1.weight"].data_ptr(), p["conv1.bias"].data_ptr(), self.a1.data_ptr(), N, s()))
        if ev_ctr is not None:
            main.wait_event(ev_ctr)
        if self.pool_fused:
            # conv2 + bias + ReLU + 2x2 max-pool + channel dropout in one kernel: the 22 MB conv2 activation never reaches memory
            ck(L.dcp_conv_flat_pool_tc(self.a1.data_ptr(), p["conv2.weight"].data_ptr(), self.pooled.data_ptr(), self.arg.data_ptr(),
                                       p["conv2.bias"].data_ptr(), N * 676, 32, 64, 9, self.off_fwd, 26, 676, 24, 24, net.dropout1.p, self.seed, 1,
                                       step_ptr, tr, 148, s()))
        else:
            ck(L.dcp_conv_flat_tc(self.a1.data_ptr(), p["conv2.weight"].data_ptr(), self.a2.data_ptr(), p["conv2.bias"].data_ptr(), None, N * 676, 32, 64, 9,
                                  self.off_fwd, 26, 676, 24, 24, 0, 148, s()))
            ck(L.dcp_convnet_pool_fwd(self.a2.data_ptr(), self.pooled.data_ptr(), self.arg.data_ptr(), N, net.dropout1.p, self.seed, 1, step_ptr, tr, s()))
        # the data-gradient's copy of conv2's filter (taps flipped, channels swapped) is taken here, while the filter is hot in L2:
        # conv2.weight may then be updated by the optimizer while the data-gradient is still running
        if not early:
            ck(tc._lib.dcp_flip_weights(p["conv2.weight"].data_ptr(), self.wt2.data_ptr(), 64, 3, 3, 32, 0, s()))
        tc.linear_fwd(self.pooled, p["fc1.weight"], p["fc1.bias"], self.h1, False)
        if ev_buf is not None:
            main.wait_event(ev_buf)
        if N <= 128 and self.fused_head:
            # BatchNorm + ReLU + dropout + fc2 + log_softmax/nll and ALL of their gradients (+ fc1's bias gradient): one 8-CTA cluster
            head_tail = [p["fc2.weight"].data_ptr(), p["fc2.bias"].data_ptr(),
                         self.loss.data_ptr(), self.logits.data_ptr(), self.dh1.data_ptr(), g["batchnorm.weight"].data_ptr(),
                         g["batchnorm.bias"].data_ptr(), g["fc2.weight"].data_ptr(), g["fc2.bias"].data_ptr(), g["fc1.bias"].data_ptr(),
                         N, bn.eps, bn.momentum, net.dropout2.p if tr else 0.0, self.seed, 2, step_ptr, tr, s()]
            if bn_push:
                # data-parallel buffer sync inside the kernel: rank 0 writes its new running statistics into every rank's block
                ck(L.dcp_convnet_head_dp(self.be.handle, self.h1.data_ptr(), self.y.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(),
                                         self.bn_flat.data_ptr(), *head_tail))
            else:
                ck(L.dcp_convnet_head(self.h1.data_ptr(), self.y.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
                                      bn.running_var.data_ptr(), bn.num_batches_tracked.data_ptr(), *head_tail))
        else:
            self._head_unfused(L, ck, s, p, g, bn, N, tr, step_ptr)
            ck(L.dcp_colsum(self.dh1.data_ptr(), g["fc1.bias"].data_ptr(), N, 128, 0, s()))
        ar_late = "ar_late" in ab
        ev_w = None
        if self.fc1_overlap:
            # fc1's weight gradient and data gradient both contract over K = 128 only: each is a short, latency-bound launch
            # that leaves the GPU half idle, and neither needs the other - so the weight gradient runs on the side stream next to
            # the data gradient (bucket 0's all-reduce and optimizer update wait for it, the convolution backward does not)
            # (128-column tiles for this pair: 72 CTAs each, so both kernels are resident at once)
            self.side.wait_stream(main)
            tc._lib.dcp_gemm_set_narrow(0)
            try:
                with torch.cuda.stream(self.side):
                    tc.gemm(self.dh1, self.pooled, g["fc1.weight"], None, 128, 9216, N, 128, 9216, 9216, 1, 1, False, 1)
                    ev_w = torch.cuda.Event(); ev_w.record(self.side)
                tc.gemm(self.dh1, p["fc1.weight"], self.dpooled, None, N, 9216, 128, 128, 9216, 9216, 0, 1, False, 1)
            finally:
                tc._lib.dcp_gemm_set_narrow(-1)
            ev0 = None
            if "b0" not in ab and self.be is not None:
                # the side stream IS the communication stream: bucket 0's all-reduce queues behind the weight gradient there
                ev0 = self._reduce_bucket(0)
        else:
            tc.gemm(self.dh1, self.pooled, g["fc1.weight"], None, 128, 9216, N, 128, 9216, 9216, 1, 1, False, 1)
            ev0 = self._reduce_bucket(0) if ("b0" not in ab and not ar_late) else None
            tc.gemm(self.dh1, p["fc1.weight"], self.dpooled, None, N, 9216, 128, 128, 9216, 9216, 0, 1, False, 1)
        if ar_late and "b0" not in ab and not self.fc1_overlap:
            ev0 = self._reduce_bucket(0)
        # fc1.weight has been read for the last time: the optimizer update of everything in bucket 0 (97 % of the
        # parameters) runs on the side stream, behind that bucket's all-reduce, underneath the convolution backward
        ev_opt0 = None
        if rows is not None:
            ent, (r0, n0, m0), _ = rows
            if ev_early is None:
                K._push_hyper(self.opt, self.opt.param_groups[0], ent)
                ck(L.dcp_optim_advance(ent["hp"].data_ptr(), s()))
            else:
                main.wait_event(ev_early)       # flipped filter + step counter: finished a forward pass ago
            opt_stream = main if "opt0main" in ab else self.side
            if "opt0end" in ab and dp_tail:
                opt_stream = None        # measurement: bucket 0's update at the very end of the main stream, between the last push and the wait
            elif opt_stream is self.side:
                self.side.wait_stream(main)
            elif ev0 is not None:
                main.wait_event(ev0); ev0 = None
            if opt_stream is main and ev_w is not None:
                main.wait_event(ev_w)
            if opt_stream is not None:
                ck(L.dcp_optim_step(ent["table"].data_ptr() + 40 * r0, n0, m0, ent["hp"].data_ptr(), self.opt._kind, 0, opt_stream.cuda_stream))
            if opt_stream is self.side:
                ev_opt0 = torch.cuda.Event(); ev_opt0.record(self.side)
        # ---- backward: convolutional trunk
        ck(L.dcp_convnet_pool_bwd(self.dpooled.data_ptr(), self.pooled.data_ptr(), self.arg.data_ptr(), self.dy2.data_ptr(), N, net.dropout1.p,
                                  self.seed, 1, step_ptr, tr, s()))
        split1, ev1a, ev1 = 0, None, None
        if self.flat_wgrad:
            ck(L.dcp_conv_wgrad_flat_tc(self.a1.data_ptr(), self.dy2.data_ptr(), self.wg_scratch.data_ptr(), N * 676, 64, 26, self.wg_parts, s()))
            if dp_tail:
                # conv2's gradients leave for every rank's payload block straight from the split-K reduction, on a second side
                # stream: the peer stores and their system-scope fences (~10 us) stay off the critical path, and nobody waits -
                # the tail kernel finds them through the signal words, this rank's own included
                ev_push = None
                ll_ok = self._ll_rows(rows[2][0], rows[2][1]) is not None and "no_ll" not in ab and "opt0end" not in ab
                pull = 0 if ("push" in ab or not ll_ok) else 1     # pull: conv2's sums stay local, the consumer reads the peers' copies
                if "push_main" in ab or (pull and "pull_main" in ab):
                    ps = main
                else:
                    ps = self.side2
                    ps.wait_stream(main)
                ck(L.dcp_splitk_reduce_push(self.be.handle, self.wg_scratch.data_ptr(), self.wg_parts, 64 * 289, self.dp_pay.data_ptr(), self.dp_slot, 0,
                                            CH_TAIL, 0, step_ptr, self.dp_ws.data_ptr(), pull, ps.cuda_stream))
                ll = self._ll_rows(rows[2][0], rows[2][1]) if ll_ok else None
                self._conv2_tail = None
                if ll is not None and pull and ps is main and "tail_early" not in ab:
                    self._conv2_tail = ll           # issued after the data-gradient (below): by then the peers' signals have long arrived
                elif ll is not None:
                    # ... followed, on the same stream, by the kernel that waits for every rank's conv2 sums, averages them and applies
                    # the optimizer update of conv2's parameters - all of it underneath the data-gradient (which reads its own copy
                    # of the filter, taken in the forward pass)
                    ent = rows[0]
                    ck(L.dcp_dp_tail(self.be.handle, ent["table"].data_ptr() + 40 * ll[1], 2, 64 * 288, ent["hp"].data_ptr(), self.opt._kind,
                                     self.buckets[1].data_ptr(), self.dp_pay.data_ptr(), self.dp_slot, CH_TAIL, self.dp_blks, 1,
                                     (C.c_int * 2)(0, 0), step_ptr, pull, ps.cuda_stream))
                if ps is not main:
                    ev_push = torch.cuda.Event(); ev_push.record(ps)
            else:
                ck(tc._lib.dcp_splitk_reduce2(self.wg_scratch.data_ptr(), g["conv2.weight"].data_ptr(), g["conv2.bias"].data_ptr(), 64 * 288,
                                              self.wg_parts, 64 * 289, s()))
                # conv2's gradients (98 % of bucket 1) are final here: reduce that part of the bucket now, under the data-gradient
                # and conv1's weight-gradient; only conv1's 320 values are left for the tail of the step
                split1 = self._bucket1_split(g) if "b1" not in ab else 0
                if split1:
                    ev1a = self.be.all_reduce_async(self.buckets[1][:split1], ReduceOp.AVG, max_ctas=self.comm_ctas)
        else:
            tc.conv_wgrad_launch(self.dy2, self.a1, g["conv2.weight"], N, 26, 26, 32, 64, 3, 3, 0, 26, 26, 8, 4, 1, 148)
            ck(L.dcp_colsum(self.dy2.data_ptr(), g["conv2.bias"].data_ptr(), N * 676, 64, 0, s()))
        ck(L.dcp_conv_flat_tc(self.dy2.data_ptr(), self.wt2.data_ptr(), self.dz1.data_ptr(), None, self.a1.data_ptr(), N * 676, 64, 32, 9,
                              self.off_bwd, 26, 676, 26, 26, 1, 148, s()))
        if dp_tail and getattr(self, "_conv2_tail", None) is not None:
            # conv2: wait for every rank's signal (sent a data-gradient ago), read the W copies of the sums - the peers' over NVLink -,
            # average, write the gradients and apply the optimizer update
            ent = rows[0]
            ck(L.dcp_dp_tail(self.be.handle, ent["table"].data_ptr() + 40 * self._conv2_tail[1], 2, 64 * 288, ent["hp"].data_ptr(), self.opt._kind,
                             self.buckets[1].data_ptr(), self.dp_pay.data_ptr(), self.dp_slot, CH_TAIL, self.dp_blks, 1,
                             (C.c_int * 2)(0, 0), step_ptr, 1, s()))
        if dp_tail:
            ent, _, (r1, n1, m1) = rows
            if self._ll_rows(r1, n1) is not None and "no_ll" not in ab and "opt0end" not in ab:
                # conv1: weight-gradient, low-latency exchange ({value, step} words, no fence / signal / kernel boundary), average and
                # optimizer update in ONE kernel - the last of the step; conv2's rows were updated on the second side stream
                ck(L.dcp_convnet_conv1_wgrad_ll(self.be.handle, self.dz1.data_ptr(), self.x.data_ptr(), N, self.c1_ll.data_ptr(),
                                                ent["table"].data_ptr() + 40 * self._ll_rows(r1, n1)[0], ent["hp"].data_ptr(), self.opt._kind,
                                                step_ptr, self.dp_ws.data_ptr() + 16, s()))
            else:
                ck(L.dcp_convnet_conv1_wgrad_push(self.be.handle, self.dz1.data_ptr(), self.x.data_ptr(), N, self.dp_pay.data_ptr(), self.dp_slot,
                                                  64 * 289, CH_TAIL, 1, step_ptr, self.dp_ws.data_ptr() + 16, s()))
                if "opt0end" in ab:
                    if ev0 is not None:
                        main.wait_event(ev0); ev0 = None
                    if ev_w is not None:
                        main.wait_event(ev_w)
                    _, (r0, n0, m0), _ = rows
                    ck(L.dcp_optim_step(ent["table"].data_ptr() + 40 * r0, n0, m0, ent["hp"].data_ptr(), self.opt._kind, 0, s()))
                # the one kernel that waits: all producers of all ranks -> average -> gradients + optimizer update of bucket 1
                ck(L.dcp_dp_tail(self.be.handle, ent["table"].data_ptr() + 40 * r1, n1, m1, ent["hp"].data_ptr(), self.opt._kind,
                                 self.buckets[1].data_ptr(), self.dp_pay.data_ptr(), self.dp_slot, CH_TAIL, self.dp_blks, 2,
                                 self._row_sig(r1, n1), step_ptr, 0, s()))
            for ev in (ev0, ev_opt0, ev_push, ev_w):   # joins AFTER the last kernel: the side streams only have to finish with the step
                if ev is not None:
                    main.wait_event(ev)
            return
        if split1 and self.fuse_tail:
            ck(L.dcp_convnet_conv1_wgrad_allreduce(self.be.handle, self.dz1.data_ptr(), self.x.data_ptr(), g["conv1.weight"].data_ptr(),
                                                   g["conv1.bias"].data_ptr(), N, self.c1_payload.data_ptr(), CH_DDP, 511, s()))
        else:
            ck(L.dcp_convnet_conv1_wgrad(self.dz1.data_ptr(), self.x.data_ptr(), g["conv1.weight"].data_ptr(), g["conv1.bias"].data_ptr(), N, s()))
            if split1:
                ev1 = self.be.all_reduce_async(self.buckets[1][split1:], ReduceOp.AVG, max_ctas=self.comm_ctas)
            elif "b1" not in ab:
                ev1 = self._reduce_bucket(1)
        if rows is not None:
            for ev in (ev1a, ev1):
                if ev is not None:
                    main.wait_event(ev)
            ent, _, (r1, n1, m1) = rows
            ck(L.dcp_optim_step(ent["table"].data_ptr() + 40 * r1, n1, m1, ent["hp"].data_ptr(), self.opt._kind, 0, s()))
            for ev in (ev0, ev_opt0, ev_w):
                if ev is not None:
                    main.wait_event(ev)
        else:
            for ev in (ev0, ev1a, ev1, ev_opt0, ev_w):
                if ev is not None:
                    main.wait_event(ev)
            self.opt.step()

    def _ll_rows(self, r1: int, n1: int):
        """(row of conv1.weight, row of conv2.weight) in the optimizer table when (conv1.weight, conv1.bias) and (conv2.weight,
        conv2.bias) are adjacent pairs there, else None."""
        params = [q for q in self.opt.param_groups[0]["params"] if q.grad is not None]
        ids = [id(params[r1 + i]) for i in range(n1)]
        net = self.net
        try:
            a, b = ids.index(id(net.conv1.weight)), ids.index(id(net.conv2.weight))
        except ValueError:
            return None
        if a + 1 >= n1 or b + 1 >= n1 or ids[a + 1] != id(net.conv1.bias) or ids[b + 1] != id(net.conv2.bias):
            return None
        return r1 + a, r1 + b

    def _row_sig(self, r1: int, n1: int):
        """For rows r1 .. r1 + n1 of the optimizer table: 0 = gradient comes from the conv2 split-K reduce, 1 = from conv1's weight-gradient."""
        params = [q for q in self.opt.param_groups[0]["params"] if q.grad is not None]
        c1 = {id(self.net.conv1.weight), id(self.net.conv1.bias)}
        return (C.c_int * n1)(*[1 if id(params[r1 + i]) in c1 else 0 for i in range(n1)])

    def _bucket1_split(self, g) -> int:
        """Number of leading elements of bucket 1 that hold conv2's gradients (0: not splittable / no communication)."""
        if self.be is None or len(self.buckets) < 2:
            return 0
        b = self.buckets[1]
        w2, b2, w1 = g["conv2.weight"], g["conv2.bias"], g["conv1.weight"]
        n2 = w2.numel() + b2.numel()
        if w2.data_ptr() != b.data_ptr() or b2.data_ptr() != b.data_ptr() + 4 * w2.numel() or w1.data_ptr() != b.data_ptr() + 4 * n2 or n2 % 4:
            return 0
        return n2

    def _opt_rows(self):
        """(cache entry, (first row, rows) of bucket 0, (first row, rows) of bucket 1) in the optimizer's device pointer
        table, or None when the table does not exist yet (first step) or the parameters of a bucket are not adjacent rows."""
        opt = self.opt
        cache = opt.__dict__.get("_dcp_cache")
        if not cache or len(opt.param_groups) != 1 or 0 not in cache or not hasattr(opt, "_kind"):
            return None
        names = dict(self.net.named_parameters())
        params = [q for q in opt.param_groups[0]["params"] if q.grad is not None]
        ent = cache[0]
        if ent["n"] != len(params) or any(q.grad.dtype != torch.float32 for q in params):
            return None
        b0 = {id(names[n]) for n in READY_ORDER[:6]}
        i0 = [i for i, q in enumerate(params) if id(q) in b0]
        i1 = [i for i, q in enumerate(params) if id(q) not in b0]
        if not i0 or not i1 or i0 != list(range(i0[0], i0[0] + len(i0))) or i1 != list(range(i1[0], i1[0] + len(i1))):
            return None
        return ent, (i0[0], len(i0), max(params[i].numel() for i in i0)), (i1[0], len(i1), max(params[i].numel() for i in i1))

    def _head_unfused(self, L, ck, s, p, g, bn, N, tr, step_ptr) -> None:
        """Kernel-per-op version of the head (any batch size); numerically the same as dcp_convnet_head."""
        net = self.net
        ck(L.dcp_bn_stats(self.h1.data_ptr(), self.bn_scratch.data_ptr(), N, 128, 0, s()))
        ck(L.dcp_bn_apply(self.h1.data_ptr(), self.h2.data_ptr(), self.bn_scratch.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(),
                          bn.running_mean.data_ptr(), bn.running_var.data_ptr(), bn.num_batches_tracked.data_ptr(), self.bn_mean.data_ptr(),
                          self.bn_rstd.data_ptr(), N, 128, float(N), bn.eps, bn.momentum, 1, 1, 0, s()))
        ck(L.dcp_dropout(self.h2.data_ptr(), self.h3.data_ptr(), N * 128, net.dropout2.p if tr else 0.0, self.seed, 2, step_ptr, 0, 1, 1, 0, s()))
        ck(L.dcp_linear_fwd_simt(self.h3.data_ptr(), p["fc2.weight"].data_ptr(), p["fc2.bias"].data_ptr(), self.logits.data_ptr(), N, 10, 128, 0, 0, s()))
        ck(L.dcp_ce_fwd(self.logits.data_ptr(), self.y.data_ptr(), self.loss.data_ptr(), self.lse.data_ptr(), N, 10, 1.0 / N, 0, s()))
        # ---- backward: head
        ck(L.dcp_ce_bwd(self.logits.data_ptr(), self.y.data_ptr(), self.lse.data_ptr(), self.one.data_ptr(), self.dlogits.data_ptr(), N, 10, 1.0 / N, 0, s()))
        ck(L.dcp_linear_wgrad_simt(self.dlogits.data_ptr(), self.h3.data_ptr(), g["fc2.weight"].data_ptr(), g["fc2.bias"].data_ptr(), N, 10, 128, 0, s()))
        ck(L.dcp_linear_dgrad_simt(self.dlogits.data_ptr(), p["fc2.weight"].data_ptr(), self.dh3.data_ptr(), N, 10, 128, 0, s()))
        ck(L.dcp_dropout(self.dh3.data_ptr(), self.dh2.data_ptr(), N * 128, net.dropout2.p if tr else 0.0, self.seed, 2, step_ptr, 0, 1, 1, 0, s()))
        ck(L.dcp_bn_bwd_reduce(self.h1.data_ptr(), self.h2.data_ptr(), self.dh2.data_ptr(), self.bn_mean.data_ptr(), self.bn_rstd.data_ptr(),
                               self.bn_scratch.data_ptr(), N, 128, 1, 0, s()))
        ck(L.dcp_bn_bwd_apply(self.h1.data_ptr(), self.h2.data_ptr(), self.dh2.data_ptr(), self.dh1.data_ptr(), self.bn_mean.data_ptr(),
                              self.bn_rstd.data_ptr(), bn.weight.data_ptr(), self.bn_scratch.data_ptr(), g["batchnorm.weight"].data_ptr(),
                              g["batchnorm.bias"].data_ptr(), N, 128, float(N), 1, 0, s()))

    def _issue_head_only(self) -> None:
        """Just the fused head kernel (used by bench/kernel_bench.py)."""
        L, net, N, bn = K.lib(), self.net, self.N, self.net.batchnorm
        p = {n: t for n, t in net.named_parameters()}
        g = {n: t.grad for n, t in p.items()}
        K._chk(L.dcp_convnet_head(self.h1.data_ptr(), self.y.data_ptr(), bn.weight.data_ptr(), bn.bias.data_ptr(), bn.running_mean.data_ptr(),
                                  bn.running_var.data_ptr(), bn.num_batches_tracked.data_ptr(), p["fc2.weight"].data_ptr(), p["fc2.bias"].data_ptr(),
                                  self.loss.data_ptr(), self.logits.data_ptr(), self.dh1.data_ptr(), g["batchnorm.weight"].data_ptr(),
                                  g["batchnorm.bias"].data_ptr(), g["fc2.weight"].data_ptr(), g["fc2.bias"].data_ptr(), g["fc1.bias"].data_ptr(),
                                  N, bn.eps, bn.momentum, net.dropout2.p, self.seed, 2, K.graph_step.data_ptr(), 1, K._stream()))

    def _reduce_bucket(self, idx: int):
        if self.be is None or idx >= len(self.buckets):
            return None
        return self.be.all_reduce_async(self.buckets[idx], ReduceOp.AVG, max_ctas=self.comm_ctas)

    def _capture(self, restore_state: bool = True) -> None:
        """Warm up (optimizer state, pointer tables, allocator) and capture the step.  The warm-up and the capture itself EXECUTE
        training steps on whatever the static input buffers hold; with ``restore_state`` the parameters, buffers and optimizer
        state are put back afterwards, so building a trainer does not train (main.py --engine fused relies on it)."""
        dev = self.dev
        opt, net = self.opt, self.net
        snap = None
        if restore_state:
            snap = {"p": [q.detach().clone() for q in net.parameters()], "b": [b.detach().clone() for b in net.buffers()],
                    "st": {q: {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in st.items()} for q, st in opt.state.items()}}
            cache = opt.__dict__.get("_dcp_cache") or {}
            snap["hp6"] = {gi: ent["hp"][6].item() for gi, ent in cache.items()}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(3):       # optimizer state, pointer tables, allocator warm-up
                self._issue()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        before = K.launch_count + (self.be.launches if self.be else 0)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._issue()
        self.launches_per_step = K.launch_count + (self.be.launches if self.be else 0) - before
        torch.cuda.synchronize(dev)
        if snap is not None:
            with torch.no_grad():
                for q, v in zip(net.parameters(), snap["p"]):
                    q.copy_(v)
                for b, v in zip(net.buffers(), snap["b"]):
                    b.copy_(v)
                for q, st in opt.state.items():
                    old = snap["st"].get(q)
                    for k, v in st.items():
                        if torch.is_tensor(v):
                            v.copy_(old[k]) if old is not None and k in old else v.zero_()     # same tensors: the captured pointers stay valid
                        elif k == "step":
                            st[k] = old[k] if old is not None and k in old else 0
                for gi, ent in (opt.__dict__.get("_dcp_cache") or {}).items():
                    ent["hp"][6] = float(snap["hp6"].get(gi, 0.0))
            if self.be is not None:
                dist.barrier()
            torch.cuda.synchronize(dev)

    def step(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """One optimisation step on batch (x, y); returns the (device, static) mean NLL loss of the batch."""
        if x.is_cuda:
            self.x.copy_(x.view(self.N, 1, 28, 28), non_blocking=True)
            self.y.copy_(y, non_blocking=True)
        else:
            self._stage_from_host(x, y)
        if self.graph is not None:
            if hasattr(self.opt, "_hyper"):
                K.sync_optimizer_hyper(self.opt)     # StepLR between replays: the captured kernels read lr from device memory
            self.graph.replay()
        else:
            self._issue()
        return self.loss

    __call__ = step

    def _stage_from_host(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Host batches go through two device staging slots filled on a copy stream: the (PCIe-speed) host->device copy
        of this step runs while the previous step is still computing; the compute stream only does a device-to-device
        copy into the buffers the captured graph reads.  Pass pinned tensors to make the copies asynchronous."""
        if self._copy_stream is None:
            f32 = dict(dtype=torch.float32, device=self.dev)
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            self._xs = [torch.empty(self.N, 1, 28, 28, **f32) for _ in range(2)]
            self._ys = [torch.empty(self.N, dtype=torch.int64, device=self.dev) for _ in range(2)]
            self._ready = [torch.cuda.Event() for _ in range(2)]
            self._free = [torch.cuda.Event() for _ in range(2)]
            self._slot = 0
        k = self._slot
        self._slot ^= 1
        main = torch.cuda.current_stream(self.dev)
        cs = self._copy_stream
        cs.wait_event(self._free[k])              # the step that last used this slot has taken its copy
        with torch.cuda.stream(cs):
            self._xs[k].copy_(x.view(self.N, 1, 28, 28), non_blocking=True)
            self._ys[k].copy_(y, non_blocking=True)
            self._ready[k].record(cs)
        main.wait_event(self._ready[k])
        self.x.copy_(self._xs[k], non_blocking=True)
        self.y.copy_(self._ys[k], non_blocking=True)
        self._free[k].record(main)
