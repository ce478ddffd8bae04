#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libvda_sm100a.so (cuobjdump -sass): the mnemonics that prove the Blackwell-native
path -- UTC*MMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG / UTMASTG / UBLKCP (TMA), UTCBAR (tcgen05.commit),
SYNCS (mbarrier), HMMA (mma.sync: the temporal-attention kernel by design), LDGSTS (cp.async), REDG / RED (the residual
epilogue).  Usage: python scripts/sass_histogram.py > profiles/sass_histogram_rNN.txt   (no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "video_depth_anything_b200", "libvda_sm100a.so")
WATCH = ["UTCHMMA", "UTCQMMA", "UTCMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCCP", "SYNCS",
         "HMMA", "LDGSTS", "LDSM", "RED", "REDG", "MUFU.EX2", "MUFU.RCP", "FFMA2", "FMUL2", "FADD2", "BAR", "LDG", "STG", "LDS", "STS"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(anonymous namespace\)::", "", name)
            name = re.sub(r"\(.*", "", name)[:120]
            cur = kernels.setdefault(name, collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_total"] += 1
            parts = op.split(".")
            cur[parts[0]] += 1
            if parts[0] == "MUFU" and len(parts) > 1:
                cur["MUFU." + parts[1]] += 1
            if parts[0].startswith("UTC") and "2CTA" in parts:
                cur[parts[0] + ".2CTA"] += 1
            if parts[0] == "UTCBAR" and "MULTICAST" in parts:
                cur["UTCBAR.MULTICAST"] += 1
    print(f"# SASS opcode histogram of {os.path.relpath(SO, ROOT)} ({os.path.getsize(SO)} bytes), sm_100a, {len(kernels)} kernels")
    print("# columns: instructions | tcgen05.mma (UTC*MMA, of which .2CTA) | tcgen05.ld/st (LDTM/STTM) | TMA (UTMALDG/UTMASTG/UBLKCP) |"
          " tcgen05.commit (UTCBAR) | mbarrier (SYNCS) | mma.sync (HMMA) | cp.async (LDGSTS) | RED/REDG | MUFU.EX2")
    tot = collections.Counter()
    for name, c in kernels.items():
        mma = sum(v for k, v in c.items() if re.fullmatch(r"UTC[A-Z]*MMA", k))
        mma2 = sum(v for k, v in c.items() if re.fullmatch(r"UTC[A-Z]*MMA\.2CTA", k))
        tma = c["UTMALDG"] + c["UTMASTG"] + c["UBLKCP"]
        row = (c["_total"], mma, mma2, c["LDTM"], c["STTM"], tma, c["UTCBAR"], c["SYNCS"], c["HMMA"], c["LDGSTS"], c["RED"] + c["REDG"],
               c["MUFU.EX2"])
        for k, v in zip("total mma mma2 ldtm sttm tma utcbar syncs hmma ldgsts red ex2".split(), row):
            tot[k] += v
        tag = "tcgen05" if mma else ("mma.sync" if c["HMMA"] else "simt")
        print(f"{row[0]:7d} | {row[1]:4d} ({row[2]:3d}) | {row[3]:4d}/{row[4]:3d} | {row[5]:4d} | {row[6]:3d} | {row[7]:4d} | {row[8]:4d} | "
              f"{row[9]:4d} | {row[10]:4d} | {row[11]:4d} | {tag:8s} {name}")
    print(f"# totals: {dict(tot)}")


if __name__ == "__main__":
    sys.exit(main())
