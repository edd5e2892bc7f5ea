#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the tcgen05 / TMEM / TMA path (no GPU needed):
    python tools/sass_counts.py > profiles/rNN_sass_tcgen05_tma_counts.txt"""
import collections, os, re, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "hip_llama.cpp_b200", "lib", "libtha_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), stdout=subprocess.PIPE, text=True).stdout.splitlines()
want = re.compile(r"\b(UTCHMMA|UTCIMMA|UTCQMMA|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|UTCBAR|UTCCP|USETMAXREG|SYNCS|REDUX|ATOMS)\b")
cnt, fn, k = collections.OrderedDict(), None, -1
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        k += 1
        fn = names[k].split("(")[0]
        cnt[fn] = collections.Counter()
        continue
    if fn:
        m = want.search(line)
        if m:
            cnt[fn][m.group(1)] += 1
print("# SASS mnemonics that prove the tcgen05 / TMEM / TMA path, per kernel of hip_llama.cpp_b200/lib/libtha_b200.so (tools/sass_counts.py)")
print("# UTCHMMA = tcgen05.mma kind::tf32 / kind::f16, UTCIMMA = kind::i8, LDTM / STTM = tcgen05.ld / st, UTMALDG = cp.async.bulk.tensor (TMA tile")
print("# load), UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit, SYNCS = mbarrier operations")
for fn in sorted(cnt):
    c = cnt[fn]
    if any(c[m] for m in ("UTCHMMA", "UTCIMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP")):
        print(f"{fn:60s} " + "  ".join(f"{m} x{c[m]}" for m in sorted(c)))
