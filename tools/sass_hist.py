"""Opcode histogram per kernel from `cuobjdump -sass libmxmoe_b200.so` (stdin or a file): the mnemonics that prove which
hardware path a kernel uses (UTCIMMA/UTCHMMA/UTCQMMA = tcgen05.mma, UTMALDG/UTMASTG/UBLKCP = TMA, LDTM/STTM = TMEM,
SYNCS = mbarrier, HMMA/IMMA = mma.sync) plus the totals.  Usage: cuobjdump -sass lib.so | c++filt | python tools/sass_hist.py"""
import collections
import re
import sys

src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
KEY = ("UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCOMMA", "UTCBAR", "UTCCP", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "LDTM", "STTM",
       "SYNCS", "HMMA", "IMMA", "REDUX", "LDGSTS", "ATOMS", "RED", "STG", "LDG", "STS", "LDS", "LDSM", "SHFL", "MUFU", "BAR", "LDL", "STL")
cur, hist = None, collections.OrderedDict()
op_re = re.compile(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Za-z0-9_]+)*)")
for line in src:
    if "Function :" in line:
        cur = line.split("Function :")[1].strip()
        hist[cur] = collections.Counter()
        continue
    m = op_re.match(line)
    if m and cur is not None:
        hist[cur][m.group(1)] += 1
        if m.group(1) in ("UTCIMMA", "UTCHMMA", "UTCQMMA", "UTMALDG", "UTMASTG", "UBLKCP", "STG", "LDG", "LDTM"):
            hist[cur][m.group(1) + m.group(2)] += 1  # keep the qualifiers (.2CTA, .128, .MULTICAST ...) for these
for name, h in hist.items():
    total = sum(v for k, v in h.items() if "." not in k)
    short = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", ""))
    print(f"== {short}   [{total} instructions]")
    keys = [k for k in KEY if h.get(k)]
    print("   " + "  ".join(f"{k}:{h[k]}" for k in keys))
    qual = sorted(k for k in h if "." in k)
    if qual:
        print("   " + "  ".join(f"{k}:{h[k]}" for k in qual))
    top = [f"{k}:{v}" for k, v in h.most_common(40) if "." not in k][:12]
    print("   top: " + "  ".join(top))
