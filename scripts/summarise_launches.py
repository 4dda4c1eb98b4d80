"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (share of total)."""
import collections, csv, re, sys

def main(path, top=30):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    tot = 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("hsfm::<unnamed>::", "").replace("void ", "")
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
        agg[name][0] += 1; agg[name][1] += v; tot += v
    print(f"| kernel | launches | total us | share | avg us |\n|---|---:|---:|---:|---:|")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"| `{k[:80]}` | {n} | {t:.1f} | {100*t/tot:.1f}% | {t/n:.2f} |")
    print(f"\ntotal {tot:.1f} us over {sum(n for n,_ in agg.values())} launches")

if __name__ == "__main__":
    main(sys.argv[1])
