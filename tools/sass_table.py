"""Count the Blackwell mnemonics per kernel in the built objects and write docs/sass/README.md + filtered listings.

    python tools/sass_table.py        (after python -m distributed_compute_pytorch_b200._build)
"""
import collections, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "build", "obj")
OUT = os.path.join(ROOT, "docs", "sass")
FILES = sorted(f[:-len(".cu.o")] for f in os.listdir(OBJ) if f.endswith(".cu.o")) if os.path.isdir(OBJ) else []   # every CUDA object of the library
COLS = [("UTC*MMA", r"UTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"UTMALDG"), ("UTMASTG / UBLKRED (TMA store / bulk reduce-add)", r"(UTMASTG|UBLKRED|UTMAREDG)"),
        ("UTCBAR", r"UTCBAR"), ("SYNCS", r"\bSYNCS"), ("packed f32x2 / FMNMX3", r"\b(FFMA2|FMUL2|FADD2|FMNMX3)"),
        ("multimem (LDGMC/STGMC/REDGMC)", r"(LDGMC|STGMC|REDGMC)"), ("PDL (ACQBULK/PREEXIT)", r"(ACQBULK|PREEXIT)"),
        ("cluster (UCGABAR/CGAERRBAR... .CLUSTER / remote smem)", r"(UCGABAR|\.CLUSTER|MAPA|CGABAR)"), ("HMMA (legacy)", r"\bHMMA")]


def main():
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for f in FILES:
        obj = os.path.join(OBJ, f + ".cu.o")
        if not os.path.exists(obj):
            continue
        sass = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
        import gzip
        with gzip.open(os.path.join(OUT, f + ".sass.gz"), "wt", compresslevel=9) as gz:      # the FULL listing of the object, next to the filtered one
            gz.write(sass)
        kern, counts = None, collections.OrderedDict()
        keep = []
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
                counts[kern] = collections.Counter()
                keep.append(line)
                continue
            if kern is None:
                continue
            hit = False
            for name, pat in COLS:
                if re.search(pat, line):
                    counts[kern][name] += 1
                    hit = True
            if hit:
                keep.append(line)
        open(os.path.join(OUT, f + ".filtered.txt"), "w").write("\n".join(keep) + "\n")
        for k, c in counts.items():
            if sum(c.values()) - c["SYNCS"] - c["PDL (ACQBULK/PREEXIT)"] <= 0 and "head" not in k and "conv1_wgrad" not in k:
                continue
            short = re.sub(r"\(anonymous namespace\)::", "", k)
            short = re.sub(r"\(.*", "", short).replace("void ", "")
            rows.append((f, short, c))
    with open(os.path.join(OUT, "README.md"), "w") as w:
        w.write("# SASS evidence (cuobjdump -sass of the sm_100a objects; regenerate with `python tools/sass_table.py`)\n\n")
        w.write("Mnemonics that prove the Blackwell-native paths (B200_PROFILING.md): `UTCHMMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG` = TMA tensor load, `UTMASTG` = TMA tensor store, `UBLKRED` = cp.reduce.async.bulk (bulk reduce-add), `FFMA2` / `FMUL2` / `FADD2` / `FMNMX3` = packed f32x2 arithmetic and three-input max, "
                "`UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier; `LDGMC` / `STGMC` / `REDGMC` = multimem.ld_reduce / multimem.red over the NVSwitch "
                "multicast mapping (multimem.st compiles to an ordinary 16-byte store whose address is the multicast VA, so it has no mnemonic of its own); `ACQBULK` / `PREEXIT` = griddepcontrol.wait / launch_dependents (programmatic dependent launch); cluster column = thread-block-cluster "
                "barrier / distributed-shared-memory mapping ops of the head kernel.  No kernel contains legacy `HMMA` (mma.sync).\n\n")
        w.write("| object | kernel | " + " | ".join(n for n, _ in COLS) + " |\n|---|---|" + "---|" * len(COLS) + "\n")
        for f, k, c in rows:
            w.write(f"| {f} | `{k[:90]}` | " + " | ".join(str(c[n]) for n, _ in COLS) + " |\n")
        w.write("\nThe `*.filtered.txt` files keep exactly the lines counted above, per object and kernel.\n")
    print(f"{len(rows)} kernels -> {OUT}/README.md")


if __name__ == "__main__":
    main()
