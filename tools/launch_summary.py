#!/usr/bin/env python
"""Per-kernel totals of ONE forward out of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`): the launches from
one `patch_gather_kernel` (first kernel of a forward without priors) up to the next one.
    python tools/launch_summary.py launches.csv [which_forward] > summary.txt"""
import collections
import csv
import re
import sys


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\((?:[^()]|\([^()]*\))*\)\s*$", "", name)  # the trailing parameter list only ("(int)-1" inside <> stays)
    name = name.replace("wm::", "")
    return name[:78]


def main():
    rows = []
    for r in csv.reader(open(sys.argv[1], errors="replace")):
        if len(r) > 10 and r[0].isdigit():
            rows.append((int(r[0]), r[4], float(r[-1]) / 1e3))
    starts = [i for i, (_, n, _) in enumerate(rows) if "patch_gather_kernel" in n]
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    if len(starts) < which + 2:
        print(f"only {len(starts)} forward starts in {len(rows)} launches", file=sys.stderr)
        sys.exit(1)
    a, b = starts[which], starts[which + 1]
    win = rows[a:b]
    tot = collections.OrderedDict()
    for _, n, t in win:
        k = short(n)
        c = tot.setdefault(k, [0, 0.0])
        c[0] += 1
        c[1] += t
    total = sum(v[1] for v in tot.values())
    print(f"one forward = launches {win[0][0]}..{win[-1][0]} of {sys.argv[1]} ({len(win)} launches, {total / 1e3:.2f} ms serialised, cold-cache):")
    print("compare SHARES with bench.py's breakdown, not absolutes.\n")
    print(f"{'kernel':80s} {'launches':>8s} {'total us':>10s} {'share':>7s}")
    for k, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:80s} {c:8d} {t:10.1f} {100 * t / total:6.1f}%")


if __name__ == "__main__":
    main()
