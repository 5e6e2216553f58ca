"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMEM / TMA use (B200_PROFILING.md):
    python profiles/sass_counts.py > profiles/r02_sass_counts.txt
UTCHMMA = tcgen05.mma (tf32 here), LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store,
UBLKCP = TMA bulk copy, LDGSTS = cp.async, UTCBAR = tcgen05.commit."""
import collections
import os
import re
import subprocess

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "sgnn_b200", "lib", "libsgnn_b200.so")
KEYS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "LDGSTS", "UTCBAR", "FFMA2", "HMMA"]

if __name__ == "__main__":
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts, fn = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            fn = re.sub(r"\(.*", "", fn)
            counts[fn] = collections.Counter()
            continue
        if fn is None:
            continue
        for k in KEYS:
            if re.search(r"\b" + k + r"\b|\b" + k + r"\.", line):
                counts[fn][k] += 1
    print("kernel".ljust(64) + "".join(k.rjust(9) for k in KEYS))
    tot = collections.Counter()
    for fn, c in counts.items():
        if sum(c.values()) == 0:
            continue
        print(fn[:63].ljust(64) + "".join(str(c[k]).rjust(9) for k in KEYS))
        tot.update(c)
    print("TOTAL".ljust(64) + "".join(str(tot[k]).rjust(9) for k in KEYS))
