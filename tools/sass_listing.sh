#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove the Blackwell-native paths (B200_PROFILING.md):
# tcgen05.mma -> UTC*MMA, tcgen05.ld/st -> LDTM/STTM, TMA bulk copies -> UBLKCP / UTMALDG, mbarriers -> SYNCS,
# packed f32 FMA -> FFMA2. usage: bash tools/sass_listing.sh > profiles/rNN_sass_listing.txt
SO=${1:-ransaclib_b200/lib/libransaclib_b200.so}
echo "# cuobjdump -sass $SO (sm_100a): occurrences per kernel"
cuobjdump -sass "$SO" 2>/dev/null | python3 -c '
import re, sys, collections
pat = re.compile(r"\b(UTC[A-Z]*MMA|LDTM|STTM|UBLKCP|UTMALDG|UTCBAR|SYNCS|FFMA2|DFMA|HMMA|ELECT|MUFU)\b")
cnt = collections.defaultdict(collections.Counter)
fn = None
for ln in sys.stdin:
    m = re.search(r"Function : (\S+)", ln)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", ln)
    if m and fn:
        mm = pat.search(m.group(1))
        if mm:
            cnt[fn][mm.group(1)] += 1
tot = collections.Counter()
for fn in sorted(cnt):
    print(fn)
    for k, v in sorted(cnt[fn].items()):
        print("    %-8s %d" % (k, v))
        tot[k] += v
print("TOTAL")
for k, v in sorted(tot.items()):
    print("    %-8s %d" % (k, v))
'
