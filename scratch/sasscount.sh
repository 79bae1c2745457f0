#!/bin/bash
# usage: sasscount.sh <mangled-substring e.g. 12BreakthroughELb0> : instruction mix of the 4-step group
# (between the "VOTE.ANY P?, P?" + "@!P BRA" that opens the group and the back-branch)
so=${2:-open_spiel-1_b200/libbsg.so}
cuobjdump -sass $so 2>/dev/null | awk -v pat="playout_kernelINS_$1" '
  /Function :/ {inf = index($0, pat) > 0}
  inf && /\/\*[0-9a-f]{4}\*\// { l=$0; sub(/\/\* 0x[0-9a-f]+ \*\/.*$/,"",l); print l }' > /tmp/sass_fn.txt
python3 - <<'PY'
import re,collections
lines=[l.rstrip() for l in open('/tmp/sass_fn.txt')]
ins=[]
for l in lines:
    m=re.match(r'\s*/\*([0-9a-f]{4})\*/\s+(.*?);',l)
    if m: ins.append((int(m.group(1),16),m.group(2).strip()))
# group = longest branch-free run
best=(0,0,0);start=0
for i,(a,t) in enumerate(ins):
    if re.search(r'\bBRA\b|\bEXIT\b|BSYNC|BSSY',t):
        if i-start>best[0]: best=(i-start,start,i)
        start=i+1
n,s,e=best
mix=collections.Counter()
for a,t in ins[s:e]:
    t=re.sub(r'^@!?U?P\d+\s+','',t)
    mix[t.split()[0]]+=1
alu=sum(v for k,v in mix.items() if re.match(r'LOP3|SHF|SEL|IADD3|ISETP|VIADD|VIMNMX|LEA|PRMT|MOV$|IABS|FLO|BREV|PLOP3|P2R|R2P',k))
fma=sum(v for k,v in mix.items() if k.startswith('IMAD'))
print("group instrs",n,"ALU",alu,"FMA(IMAD)",fma,"POPC",mix.get('POPC',0),"LDS",sum(v for k,v in mix.items() if k.startswith('LDS')))
print(sorted(mix.items(),key=lambda kv:-kv[1])[:14])
PY
