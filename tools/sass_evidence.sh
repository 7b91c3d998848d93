#!/bin/sh
# Per-kernel counts of the SASS mnemonics that prove which hardware paths a kernel uses
# (tcgen05 / TMEM / TMA / mbarrier / cluster barriers / fp64).  usage: sh tools/sass_evidence.sh [lib.so]
LIB=${1:-pilottai_b200/libpilott_b200.so}
cuobjdump -sass "$LIB" 2>/dev/null | python3 -c "
import sys, re, collections, subprocess
fn=None; c=collections.defaultdict(collections.Counter)
for line in sys.stdin:
    m=re.search(r'Function : (\S+)', line)
    if m: fn=m.group(1); continue
    m=re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)', line)
    if m and fn: c[fn][m.group(1)]+=1
keys=['UTCHMMA','UTCQMMA','UTCBAR','UTMALDG','UTMAPF','UBLKCP','LDTM','SYNCS','UCGABAR_ARV','UCGABAR_WAIT','UTCATOMSWS','DFMA','DADD','POPC','LDGSTS','REDG','ATOMG','MEMBAR','FMNMX']
for f,cc in c.items():
    d=subprocess.run(['c++filt',f],capture_output=True,text=True).stdout.strip().split('(')[0]
    print(f'{d:60s} total {sum(cc.values()):6d}  ' + '  '.join(f'{k} {cc[k]}' for k in keys if cc[k]))
" | sed 's/pilott:://' | sort
