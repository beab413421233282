"""The reference program on GPUs: main.py (DDP over the NVLink backend, hand-written kernels, mnist.pt with module.*
keys), the cbasics primitive demo, and the bounded-wait failure path (a missing peer becomes an error, not a hang)."""
import os
import subprocess
import sys
import time

import pytest
import torch

import distributed_compute_pytorch_b200 as dcp
from tests.helpers import free_port, init

pytestmark = [pytest.mark.gpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.multigpu
def test_main_py_two_gpus(tmp_path):
    env = dict(os.environ, MASTER_PORT=str(free_port()), MASTER_ADDR="127.0.0.1", PYTHONPATH=ROOT)
    ck = tmp_path / "mnist.pt"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "--epochs", "2", "--batch_size", "64", "--data-size", "2048", "--gpus", "2",
                        "--checkpoint", str(ck), "--lr", "1.0"], capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = r.stdout
    assert "running DDP example on rank 0" in out and "running DDP example on rank 1" in out
    assert "epoch: 0 [0/16" in out and "Test set: Average loss:" in out and out.count("time to complete this epoch:") == 4
    sd = torch.load(ck, map_location="cpu")
    assert list(sd.keys())[0] == "module.conv1.weight" and tuple(sd["module.fc1.weight"].shape) == (128, 9216)
    assert sd["module.batchnorm.num_batches_tracked"].item() == 32
    # loss printed at the start of epoch 1 is lower than at the start of epoch 0 (it learns on the synthetic classes)
    losses = [float(l.split("Loss:")[1]) for l in out.splitlines() if "Loss:" in l and "[0/16" in l]
    assert len(losses) == 2 and losses[1] < losses[0], losses


@pytest.mark.multigpu
def test_cbasics_demo_two_gpus():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "demos", "cbasics.py"), "--gpus", "2"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "cbasics: all primitives OK" in r.stdout


def _absent_peer(rank, world, port):
    os.environ["DCP_COLL_TIMEOUT_S"] = "3"
    torch.cuda.set_device(rank)
    dist = init(rank, world, port, "nccl")
    t = torch.ones(1024, device=f"cuda:{rank}")
    if rank == 0:
        t0 = time.time()
        try:
            dist.all_reduce(t)              # rank 1 never joins this collective
            torch.cuda.synchronize()
            dist.distributed._world().cuda.check_health()
            raise AssertionError("expected the bounded wait to report a missing peer")
        except RuntimeError as e:
            assert "peer" in str(e) or "timed out" in str(e), str(e)
        assert time.time() - t0 < 30
    else:
        time.sleep(8)
    os._exit(0)     # the group is deliberately left inconsistent: skip the collective teardown


@pytest.mark.multigpu
def test_missing_peer_is_an_error_not_a_hang():
    dcp.spawn(_absent_peer, args=(2, free_port()), nprocs=2, join=True)


@pytest.mark.gpu
def test_main_py_engines_agree_on_one_gpu(tmp_path):
    """python main.py on ONE GPU with --engine eager (layer-by-layer autograd, the reference's loop body) and --engine fused (the
    hand-scheduled flagship step) from the same seed, dropout off: same prints, and checkpoints that agree.  1000 samples at
    batch 128 = 7 full batches through the engine + a last batch of 104 that takes the eager path inside the same pass.
    Adadelta's first updates are sign-like (|delta| ~ sqrt(eps / (1 - rho)) wherever a gradient is tiny), so agreement is
    asserted statistically: almost every weight within 2e-3 after 16 updates at lr = 1, running statistics close on average."""
    sds, outs = {}, {}
    for eng in ("eager", "fused"):
        env = dict(os.environ, MASTER_PORT=str(free_port()), MASTER_ADDR="127.0.0.1", PYTHONPATH=ROOT)
        ck = tmp_path / f"mnist_{eng}.pt"
        r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "--epochs", "2", "--batch_size", "128", "--data-size", "1000", "--gpus", "1",
                            "--checkpoint", str(ck), "--lr", "1.0", "--engine", eng, "--no-dropout"], capture_output=True, text=True, timeout=600,
                           env=env, cwd=str(tmp_path))
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
        outs[eng] = r.stdout
        sds[eng] = torch.load(ck, map_location="cpu")
        assert os.path.exists(str(ck) + ".state")
    for eng, out in outs.items():
        assert "epoch: 0 [0/8" in out and "epoch: 1 [0/8" in out and "Test set: Average loss:" in out, (eng, out[-1500:])
    a, b = sds["eager"], sds["fused"]
    assert list(a.keys()) == list(b.keys()) and list(a.keys())[0] == "conv1.weight"
    assert a["batchnorm.num_batches_tracked"].item() == b["batchnorm.num_batches_tracked"].item() == 16
    for k in a:
        if a[k].dtype != torch.float32:
            continue
        d = (a[k] - b[k]).abs()
        if "running" in k:
            assert d.mean().item() < 0.02 and d.max().item() < 0.15, (k, d.mean().item(), d.max().item())     # statistics of slightly different weights
        else:
            assert (d > 2e-3).float().mean().item() < 0.02, (k, d.max().item(), (d > 2e-3).float().mean().item())
    first = {e: float([l for l in o.splitlines() if "Loss:" in l][0].split("Loss:")[1]) for e, o in outs.items()}
    assert abs(first["eager"] - first["fused"]) < 1e-3, first
