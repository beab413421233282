"""Install the UNMODIFIED reference into baseline/_ref (git-ignored; travels to the GPU box).

The prescribed command is tried first::

    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref /root/reference

It cannot succeed: the reference is three loose files (main.py, cbasics.sh, requirements.txt) with no
setup.py / pyproject.toml, so pip answers "Directory '/root/reference' is not installable" (also from a
/tmp copy with --no-deps; outcome recorded in DESIGN.md).  The fallback is the only install such a
script has: its files are copied byte-for-byte and the sha256 of each copy is checked against the
source.  bench.py --impl reference then imports baseline/_ref/main.py as a module and drives the
reference's own ConvNet / train() through torch DDP + NCCL.
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = "/root/reference"


def sha(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def install(verbose=True) -> bool:
    if os.path.exists(os.path.join(DST, "main.py")):
        return True
    if not os.path.isdir(SRC):
        return False
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", "/opt/wheelhouse",
                        "--target", DST, SRC], capture_output=True, text=True)
    if verbose:
        print("pip:", (r.stdout + r.stderr).strip().splitlines()[-1] if (r.stdout + r.stderr).strip() else r.returncode)
    if r.returncode == 0 and os.path.exists(os.path.join(DST, "main.py")):
        return True
    os.makedirs(DST, exist_ok=True)
    for name in sorted(os.listdir(SRC)):
        s, d = os.path.join(SRC, name), os.path.join(DST, name)
        if os.path.isfile(s):
            shutil.copyfile(s, d)
            assert sha(s) == sha(d), name
            if verbose:
                print(f"copied {name} sha256={sha(d)[:16]}")
    return os.path.exists(os.path.join(DST, "main.py"))


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
