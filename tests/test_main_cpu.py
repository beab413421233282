"""BASELINE.json configs[0]: main.py small CNN DDP on synthetic MNIST-shaped data, world_size=2 on
CPU - runs to completion, ranks hold identical params, mnist.pt keys match the reference's."""
import os
import subprocess
import sys

import torch

import main as app

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cli_flags_match_reference():
    opt = app.build_parser().parse_args([])
    assert (opt.batch_size, opt.lr, opt.epochs, opt.gamma, opt.gpus, opt.no_cuda) == (128, 0.001, 20, 0.7, 4, False)
    assert app.build_parser().parse_args(["--no-cuda", "True"]).no_cuda is True
    assert app.build_parser().parse_args(["--no-cuda", "False"]).no_cuda is False  # the reference treats 'False' as truthy


def test_main_world2_cpu(tmp_path, free_port):
    env = dict(os.environ, MASTER_PORT=str(free_port), MASTER_ADDR="127.0.0.1", PYTHONPATH=ROOT)
    ck = tmp_path / "mnist.pt"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "--epochs", "2", "--batch_size", "16", "--data-size", "256",
                        "--no-cuda", "True", "--checkpoint", str(ck), "--lr", "1.0"],
                       capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout + r.stderr
    out = r.stdout
    assert "running DDP example on rank 0" in out and "running DDP example on rank 1" in out
    assert "epoch: 0 [0/8" in out and "Loss:" in out
    assert "Test set: Average loss:" in out and "Accuracy:" in out
    assert out.count("time to complete this epoch:") == 4  # every rank, every epoch
    sd = torch.load(ck)
    expect = ["module." + k for k in ("conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias",
                                      "fc2.weight", "fc2.bias", "batchnorm.weight", "batchnorm.bias", "batchnorm.running_mean",
                                      "batchnorm.running_var", "batchnorm.num_batches_tracked")]
    assert list(sd.keys()) == expect
    assert tuple(sd["module.fc1.weight"].shape) == (128, 9216) and sd["module.batchnorm.num_batches_tracked"].item() == 16
    assert hasattr(sd, "_metadata")


def test_tracer_cpu(tmp_path):
    """utils.Tracer: ranges, summary and the chrome-trace export work without a GPU (host row only)."""
    import json
    import time
    from distributed_compute_pytorch_b200.utils import Tracer
    tr = Tracer(enabled=True, rank=3)
    with tr.range("a"):
        time.sleep(0.01)
    with tr.range("a"):
        pass
    tr.mark("m")
    s = tr.summary()
    assert s["a"]["count"] == 2 and s["a"]["host_ms"] >= 9.0
    path = tr.export_chrome(str(tmp_path / "t.json"))
    ev = json.load(open(path))["traceEvents"]
    assert any(e.get("name") == "a" and e.get("ph") == "X" and e["pid"] == 3 for e in ev)
    off = Tracer(enabled=False)
    with off.range("x"):
        pass
    assert off.summary() == {}


def test_checkpoint_helpers_roundtrip(tmp_path):
    """utils.checkpoint: reference-format model file (module.* prefix handling both ways) + training-state sidecar."""
    import torch
    import distributed_compute_pytorch_b200 as dcp
    from distributed_compute_pytorch_b200 import ops
    from distributed_compute_pytorch_b200.utils import save_model, load_model, save_training_state, load_training_state
    torch.manual_seed(0)
    m = dcp.models.ConvNet()
    opt = ops.Adadelta(m.parameters(), lr=0.5); sch = ops.StepLR(opt, 1, 0.7)
    x = torch.randn(4, 1, 28, 28); y = torch.randint(0, 10, (4,))
    ops.nll_loss(m(x), y).backward(); opt.step(); sch.step()
    path = str(tmp_path / "mnist.pt")
    save_model(path, m); save_training_state(path + ".state", opt, sch, epoch=1)
    sd = torch.load(path)
    assert list(sd.keys())[0] == "conv1.weight" and len(sd) == 13 and tuple(sd["fc1.weight"].shape) == (128, 9216)
    # load into a "DDP-shaped" wrapper (module.* keys) and back into a plain model
    class Wrap(torch.nn.Module):
        def __init__(self, mod):
            super().__init__(); self.module = mod
    w = Wrap(dcp.models.ConvNet()); load_model(path, w)
    assert torch.equal(w.module.fc2.weight, m.fc2.weight)
    torch.save(w.state_dict(), path)                       # module.* keys, as the reference writes under DDP
    m2 = dcp.models.ConvNet(); load_model(path, m2)
    assert torch.equal(m2.conv2.weight, m.conv2.weight)
    opt2 = ops.Adadelta(m2.parameters(), lr=0.5); sch2 = ops.StepLR(opt2, 1, 0.7)
    st = load_training_state(path + ".state", opt2, sch2)
    assert st["epoch"] == 1 and sch2.last_epoch == sch.last_epoch
    a = opt.state_dict()["state"][0]["square_avg"]; b = opt2.state_dict()["state"][0]["square_avg"]
    assert torch.equal(a, b)


def test_resume_continues_epoch_and_lr(tmp_path, free_port):
    """A first run always leaves the optimizer / schedule / epoch sidecar next to the checkpoint (written before the model
    file); --resume picks both up: the second run starts at the stored epoch with the decayed learning rate."""
    ck = tmp_path / "mnist.pt"
    base = [sys.executable, os.path.join(ROOT, "main.py"), "--batch_size", "16", "--data-size", "64", "--no-cuda", "True",
            "--checkpoint", str(ck), "--lr", "1.0", "--gamma", "0.5"]
    env = dict(os.environ, MASTER_PORT=str(free_port), MASTER_ADDR="127.0.0.1", PYTHONPATH=ROOT)
    r = subprocess.run(base + ["--epochs", "1"], capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout + r.stderr
    state = torch.load(str(ck) + ".state", weights_only=False)
    assert state["epoch"] == 1 and abs(state["optimizer"]["param_groups"][0]["lr"] - 0.5) < 1e-12
    r = subprocess.run(base + ["--epochs", "3", "--resume"], capture_output=True, text=True, timeout=600, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "resumed from" in r.stdout and "at epoch 1" in r.stdout
    assert "epoch: 1 [0/2" in r.stdout and "epoch: 2 [0/2" in r.stdout and "epoch: 0 [" not in r.stdout
    state = torch.load(str(ck) + ".state", weights_only=False)
    assert state["epoch"] == 3 and abs(state["optimizer"]["param_groups"][0]["lr"] - 0.125) < 1e-12
