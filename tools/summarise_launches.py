"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel shares of ONE decoded token.

    python tools/summarise_launches.py gpurun_out/launches.csv <launches_per_token> [token_index] > profiles/rNN_launches_*.txt

The capture covers several tokens of `bench.py --steps 2 --warmup 3`; the token_index-th complete token (default: the last one
that fits) is cut out by position.  ncu times are cold-cache and serialised: compare SHARES, not absolutes."""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path, per_token = sys.argv[1], int(sys.argv[2])
    rows = []
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r["Metric Name"] == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], float(r["Metric Value"]) / 1e3, r["Grid Size"], r["Block Size"]))
    n_tok = len(rows) // per_token
    idx = int(sys.argv[3]) if len(sys.argv) > 3 else n_tok - 1
    # the first launches of the run are the same token sequence: tokens start at multiples of per_token
    cut = rows[idx * per_token:(idx + 1) * per_token]
    agg = OrderedDict()
    for name, us, grid, block in cut:
        key = re.sub(r"\(.*", "", name).replace("void ", "").replace("b200::", "")
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += us
    total = sum(a[1] for a in agg.values())
    print(f"# {len(rows)} launches captured, {per_token} per token; token {idx} (launches {idx * per_token}..{(idx + 1) * per_token - 1})")
    for key, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{key:58s} {n:4d} launches {us:9.1f} us  avg {us / n:7.2f} us  share {100 * us / total:5.1f}%")
    print(f"total {total:.1f} us over {len(cut)} launches (ncu per-launch times are cold-cache and serialised: compare shares)")


if __name__ == "__main__":
    main()
