"""Turn a bench/coll_sweep.py JSONL into the markdown tables kept under profiles/.

    python tools/sweep_to_md.py profiles/coll_sweep_8gpu_run2.jsonl "8 x B200, run 2" > profiles/coll_sweep_8gpu_run2.md
"""
import collections
import json
import sys


def main():
    path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    rows = [json.loads(l) for l in open(path) if l.strip()]
    by = collections.defaultdict(dict)
    impls = collections.defaultdict(list)
    for r in rows:
        by[(r["op"], r["bytes"])][r["impl"]] = r
        if r["impl"] not in impls[r["op"]]:
            impls[r["op"]].append(r["impl"])
    print(f"# Collective sweep, {title} (device-timed, CUDA events, max over ranks; `us / bus GB/s`)\n")
    print("`*_sym` = tensors inside the symmetric heap (zero copy); `ours` / `*_staged` / `auto` = ordinary torch tensors bounced through the "
          "staging area inside the same kernel; `nccl` = torch.distributed NCCL on the same box.  **bold** = fastest of the row.\n")
    for op in sorted(impls):
        cols = [c for c in impls[op] if c != "nccl"] + (["nccl"] if "nccl" in impls[op] else [])
        print(f"\n## {op}\n")
        print("| bytes | " + " | ".join(cols) + " | nccl / best ours | nccl / public API (ordinary tensor) | SM MHz |")
        print("|---|" + "---|" * (len(cols) + 3))
        for (o, nbytes) in sorted(k for k in by if k[0] == op):
            cell = by[(o, nbytes)]
            best = min((cell[c]["us"] for c in cols if c in cell), default=None)
            out = []
            for c in cols:
                if c not in cell:
                    out.append("-")
                    continue
                txt = f'{cell[c]["us"]:.1f} / {cell[c]["bus_GBps"]:.0f}'
                out.append(f"**{txt}**" if cell[c]["us"] == best else txt)
            ours = min((cell[c]["us"] for c in cols if c in cell and c != "nccl"), default=None)
            ratio = f'{cell["nccl"]["us"] / ours:.2f}x' if ours and "nccl" in cell else "-"
            pub = cell.get("auto") or cell.get("ours")        # what dist.<collective>(ordinary tensor) runs
            pratio = f'{cell["nccl"]["us"] / pub["us"]:.2f}x' if pub and "nccl" in cell else "-"
            mhz = next((str(v["clocks"].get("sm_mhz")) for v in cell.values() if isinstance(v.get("clocks"), dict)), "-")
            print(f"| {nbytes} | " + " | ".join(out) + f" | {ratio} | {pratio} | {mhz} |")


if __name__ == "__main__":
    main()
