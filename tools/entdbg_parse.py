#!/usr/bin/env python3
"""Per-SM load and chain times of k_entropy_split from the printf lines of an ENT_DEBUG build
(tools/ab_build.sh entdbg "-DENT_DEBUG"; WEBPB200_SO=ab/libwebpb200_entdbg.so python tools/entropy_ab.py 1024 > log):
usage: tools/entdbg_parse.py log"""
import sys, collections
rows=[l.split() for l in open(sys.argv[1]) if l.startswith('ENT ')]
# keep the last launch: take last 2048 lines
rows=rows[-2048:]
tok=[(int(r[2]),int(r[3]),int(r[5]),int(r[6]),int(r[4])) for r in rows if r[1]=="0"]
mod=[(int(r[2]),int(r[3]),int(r[5]),int(r[6])) for r in rows if r[1]=="1"]
t0=min(x[2] for x in tok+[(m[0],m[1],m[2],m[3],0) for m in mod])
persm=collections.Counter(x[1] for x in tok)
print('token chains per SM histogram:', sorted(collections.Counter(persm.values()).items()))
import statistics
for k in sorted(set(persm.values())):
    d=[(x[3]-x[2])/1e6 for x in tok if persm[x[1]]==k]
    print('SMs with %d chains: chain ms mean %.2f max %.2f' % (k, statistics.mean(d), max(d)))
print('token start (ms) min/max: %.2f %.2f ; end max %.2f' % (min((x[2]-t0)/1e6 for x in tok), max((x[2]-t0)/1e6 for x in tok), max((x[3]-t0)/1e6 for x in tok)))
print('mode chains: dur mean %.2f max %.2f' % (statistics.mean((m[3]-m[2])/1e6 for m in mod), max((m[3]-m[2])/1e6 for m in mod)))
ws=collections.Counter((x[1], x[4] % 4) for x in tok)
print('token chains per (SM, warpid%4) histogram:', sorted(collections.Counter(ws.values()).items()))
