"""Static evidence from the built library, no GPU needed: per kernel the registers / stack / static shared memory
(`cuobjdump -res-usage`) and the count of the SASS instructions that prove the Blackwell paths — UTCHMMA / UTCQMMA
(tcgen05.mma), UTMALDG / UTMASTG / UTMAREDG (TMA load / store / reduce-add), UBLKCP (cp.async.bulk, non-tensor), UTCBAR (tcgen05.commit), LDTM / STTM (tcgen05.ld /
st), SYNCS (mbarrier) — from `cuobjdump -sass`.  Usage: python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "hunyuanworld-mirror_b200", "lib", "libwm_b200.so")
MNEMONICS = ("UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UTCBAR", "LDTM", "STTM", "SYNCS", "MUFU.EX2", "HMMA", "STL", "LDL")


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(.*", "", n).replace("void ", "").replace("wm::", "") for n in out]


def main():
    if not os.path.exists(SO):
        sys.exit(f"{SO} is missing: run python __graft_entry__.py first")
    res = subprocess.run(["cuobjdump", "-res-usage", SO], capture_output=True, text=True).stdout.splitlines()
    usage, cur = {}, None
    for line in res:
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line:
            usage[cur] = dict(kv.split(":") for kv in line.split() if ":" in kv and not kv.startswith("CONSTANT"))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout.splitlines()
    counts, cur = collections.defaultdict(collections.Counter), None
    for line in sass:
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            continue
        if cur:
            for mn in MNEMONICS:
                if re.search(r"\b" + re.escape(mn), line):
                    counts[cur][mn] += 1
    names = sorted(usage)
    pretty = dict(zip(names, demangle(names)))
    print(f"{os.path.relpath(SO, ROOT)}: {len(names)} kernels, sm_100a (cuobjdump -res-usage / -sass)\n")
    print(f"{'kernel':58s} {'regs':>4s} {'stack':>5s} {'smem':>5s}  " + " ".join(f"{m:>8s}" for m in MNEMONICS))
    for n in sorted(names, key=lambda k: pretty[k]):
        u, c = usage[n], counts.get(n, {})
        print(f"{pretty[n][:58]:58s} {u.get('REG', '?'):>4s} {u.get('STACK', '?'):>5s} {u.get('SHARED', '?'):>5s}  " +
              " ".join(f"{c.get(m, 0):8d}" for m in MNEMONICS))


if __name__ == "__main__":
    main()
