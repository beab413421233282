"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (second half of the file = the last step).

    python tools/launches_to_md.py gpurun_out/launches_gpt2.csv [top_n]
"""
import collections, csv, re, sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    hdr, data = rows[hi], rows[hi + 1:]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    data = data[len(data) // 2:]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in data:
        name = re.sub(r"\(.*", "", r[kn]).replace("void ", "").replace("<unnamed>::", "")[:80]
        v = float(r[mv].replace(",", ""))
        v = v / 1000 if r[mu] == "ns" else (v * 1000 if r[mu] == "ms" else v)
        agg[name][0] += 1; agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"| kernel | launches | total us | share |\n|---|---|---|---|")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"| `{k}` | {v[0]} | {v[1]:.0f} | {100 * v[1] / tot:.1f} % |")
    print(f"| all {sum(v[0] for v in agg.values())} launches | | {tot:.0f} | (cold-cache, serialised under ncu: compare shares) |")


if __name__ == "__main__":
    main()
