"""Per-kernel counts of the SASS mnemonics that prove the Blackwell data path of libssf_b200.so
(B200_PROFILING.md): UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UBLKCP (cp.async.bulk),
UTMALDG (tensor-map TMA load), UTCBAR (tcgen05.commit), SYNCS (mbarrier), plus registers per thread.

    python scripts/sass_summary.py > profiles/sass_summary.txt
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ska_sdp_screen_fitting_b200", "libssf_b200.so")
MNEMONICS = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "MUFU", "DFMA", "STG", "PRMT"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout
    return out.splitlines()


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    counts = {}
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = dict.fromkeys(MNEMONICS, 0)
            continue
        if cur is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            op = m.group(1).split(".")[0]
            if op in counts[cur]:
                counts[cur][op] += 1
    names = list(counts)
    nice = demangle(names)
    print(f"# SASS summary of {os.path.relpath(LIB, ROOT)} (cuobjdump -sass / -res-usage), sm_100a")
    print(f"{'kernel':<70}{'regs':>5}{'smem':>7}" + "".join(f"{m:>9}" for m in MNEMONICS))
    for raw, name in sorted(zip(names, nice), key=lambda p: p[1]):
        short = re.sub(r"\(anonymous namespace\)::", "", name)
        short = re.sub(r"\(.*", "", short)
        short = re.sub(r"^void ", "", short)
        r, s = regs.get(raw, (0, 0))
        print(f"{short[:69]:<70}{r:>5}{s:>7}" + "".join(f"{counts[raw][m]:>9}" for m in MNEMONICS))
    tot = {m: sum(c[m] for c in counts.values()) for m in MNEMONICS}
    print(f"{'TOTAL':<82}" + "".join(f"{tot[m]:>9}" for m in MNEMONICS))


if __name__ == "__main__":
    sys.exit(main())
