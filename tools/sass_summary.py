"""Per-kernel counts of the Blackwell-specific SASS mnemonics in singa_b200/libefl.so (cuobjdump -sass): UTCHMMA
(tcgen05.mma), UTMALDG / UTMASTG (TMA tensor loads / stores), LDTM (tcgen05.ld), UTCBAR (tcgen05.commit), UBLKCP (bulk
copies), SYNCS (mbarrier).  Evidence that the hot kernels are tcgen05 / TMA code, committed under profiles/.

  python tools/sass_summary.py > profiles/r2_sass_mnemonics.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "singa_b200", "libefl.so")
MNEMONICS = ("UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR", "UBLKCP", "SYNCS", "HMMA", "FFMA", "MUFU")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    counts, order, cur, k = collections.defaultdict(collections.Counter), [], None, 0
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = names[k]
            k += 1
            order.append(cur)
            continue
        if cur is None:
            continue
        m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            op = m.group(1)
            counts[cur]["_total"] += 1
            for mn in MNEMONICS:
                if op.startswith(mn):
                    counts[cur][mn] += 1
    sha = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# SASS mnemonic counts per kernel, singa_b200/libefl.so built from git {sha} (cuobjdump -sass, sm_100a)")
    print("# kernel | instructions | " + " | ".join(MNEMONICS))
    tot = collections.Counter()
    for name in order:
        c = counts[name]
        short = re.sub(r"\(.*", "", name)
        if not any(c[m] for m in MNEMONICS[:7]) and "--all" not in sys.argv:
            continue
        print(f"{short} | {c['_total']} | " + " | ".join(str(c[m]) for m in MNEMONICS))
        tot.update(c)
    print("# library total of the listed kernels | " + str(tot["_total"]) + " | " + " | ".join(str(tot[m]) for m in MNEMONICS))


if __name__ == "__main__":
    main()
