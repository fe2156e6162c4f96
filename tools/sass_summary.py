"""Per-kernel counts of the SASS mnemonics that prove the Blackwell paths (tcgen05 MMA, TMEM loads, TMA, async copies) in the
built library: python tools/sass_summary.py > profiles/sass_summary.txt   (runs cuobjdump -sass, no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fcaf3d_b200", "libfcaf3d_b200.so")
PAT = collections.OrderedDict([("UTCHMMA", r"\bUTCHMMA"), ("UTCQMMA/other UTC*MMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"),
                               ("UTMALDG", r"\bUTMALDG"), ("UTCBAR", r"\bUTCBAR"), ("LDGSTS", r"\bLDGSTS"),
                               ("SYNCS", r"\bSYNCS"), ("ATOM/RED", r"\b(ATOM|ATOMG|RED|REDG)\b"), ("MATCH", r"\bMATCH")])


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        if cur is None:
            continue
        for name, pat in PAT.items():
            if re.search(pat, line):
                cur[name] += 1
        cur["instructions"] += 1 if re.search(r"/\*[0-9a-f]{4}\*/", line) else 0
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    print("# cuobjdump -sass fcaf3d_b200/libfcaf3d_b200.so (sm_100a): mnemonic counts per kernel; only kernels that use")
    print("# tensor cores, TMEM, TMA or cp.async are listed, then library totals")
    tot = collections.Counter()
    for (mangled, c), name in zip(kernels.items(), demangle):
        tot.update(c)
        if any(c[k] for k in ("UTCHMMA", "UTCQMMA/other UTC*MMA", "LDTM", "UTMALDG", "LDGSTS")):
            short = re.sub(r"\(.*", "", name)[:90]
            print("%-92s %s" % (short, " ".join("%s=%d" % (k, c[k]) for k in PAT if c[k])))
    print("TOTAL kernels=%d %s" % (len(kernels), " ".join("%s=%d" % (k, tot[k]) for k in PAT)))


if __name__ == "__main__":
    main()
