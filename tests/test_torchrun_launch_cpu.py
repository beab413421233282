"""The launch mode the benchmark driver uses: ``python -m torch.distributed.run --master-addr A --master-port P``
(static rendezvous).  The launcher keeps its own store on MASTER_PORT (TORCHELASTIC_USE_AGENT_STORE=True), so
``init_process_group`` must open this package's store elsewhere and publish the port through the launcher's store
(store.py::_rendezvous_behind_agent).  CPU / gloo, world size 2, including a destroy + second init."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_static_torchrun_rendezvous(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(textwrap.dedent(f"""
        import sys, torch
        sys.path.insert(0, {ROOT!r})
        import distributed_compute_pytorch_b200 as dcp
        for attempt in range(2):
            dcp.init_process_group("gloo")
            t = torch.ones(4) * (dcp.get_rank() + 1)
            dcp.all_reduce(t)
            assert t.tolist() == [3.0] * 4, t
            print("OK", attempt, dcp.get_rank(), dcp.get_world_size(), flush=True)
            dcp.destroy_process_group()
    """))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(script)], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("OK") == 4, r.stdout + r.stderr
