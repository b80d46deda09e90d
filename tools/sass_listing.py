"""profiles/*_sass_tcgen05.txt: per kernel of libdroid_b200.so the counts of the tensor-core / TMEM / TMA SASS mnemonics
(UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA load / store, UTCBAR = tcgen05.commit, SYNCS =
mbarrier) and the first line of each.  Usage: python tools/sass_listing.py out.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "droid-slam_b200", "lib", "libdroid_b200.so")
KEYS = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "UTCATOMSWS", "SYNCS", "UTMAPF", "UTCCP")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    lines = ["cuobjdump -sass droid-slam_b200/lib/libdroid_b200.so (sm_100a): tensor-core / TMEM / TMA mnemonics per kernel", ""]
    name, cnt, first = None, None, None

    def flush():
        if name and any(cnt[k] for k in ("UTCHMMA", "LDTM", "UTMALDG", "UTMASTG")):
            lines.append("== " + name)
            lines.append("   " + str({k: cnt[k] for k in KEYS if cnt[k]}))
            for k in KEYS:
                if k in first:
                    lines.append("   " + first[k].strip())
            lines.append("")
    for ln in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            flush()
            name, cnt, first = m.group(1), collections.Counter(), {}
            continue
        if name:
            for k in KEYS:
                if re.search(r"\b%s\b" % k, ln):
                    cnt[k] += 1
                    first.setdefault(k, ln)
    flush()
    open(sys.argv[1], "w").write("\n".join(lines) + "\n")
    print("wrote", sys.argv[1], len(lines), "lines")


if __name__ == "__main__":
    main()
