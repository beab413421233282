import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cbasics_demo_on_cpu_mesh():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "demos", "cbasics.py"), "--gpus", "3", "--cpu"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "cbasics: all primitives OK" in r.stdout and "reduce_scatter" in r.stdout
