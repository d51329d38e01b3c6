#!/bin/bash
# usage: sass_lines.sh <kernel-substring> [N]  — SASS instruction count per source line of a kernel in liblgasr.so
cd "$(dirname "$0")/../llm_guided_asr_b200/csrc" || exit 1
tmp=$(mktemp -d); cp liblgasr.so $tmp/; cd $tmp
cuobjdump -xelf all liblgasr.so >/dev/null 2>&1
nvdisasm -g *.cubin 2>/dev/null | python3 -c "
import sys,re,collections
pat=sys.argv[1]; N=int(sys.argv[2])
cur=None; cnt=collections.Counter(); on=False
for l in sys.stdin:
    if l.startswith('//---') and '.text.' in l: on = pat in l; continue
    if not on: continue
    m=re.search(r'//## File \"([^\"]+)\", line (\d+)',l)
    if m: cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/',l): cnt[cur]+=1
for k,v in sorted(cnt.items(), key=lambda x:-x[1])[:N]: print(v,k)
print('total',sum(cnt.values()))
" "$1" "${2:-30}"
rm -rf $tmp
