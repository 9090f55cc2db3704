#!/usr/bin/env python
"""Count SASS opcodes per kernel in a cubin/.so (dev tool): python tools_sass_count.py file [filter]"""
import re, collections, subprocess, sys
txt = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
flt = sys.argv[2] if len(sys.argv) > 2 else ""
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n')[0]
    if flt and flt not in name:
        continue
    m = re.search(r'ILi(\d+)E', name)
    V = int(m.group(1)) if m else 1
    ops = collections.Counter()
    for line in f.split('\n'):
        mm = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)', line)
        if mm:
            op = mm.group(2)
            ops[op if op.startswith('IMAD') else op.split('.')[0]] += 1
    tot = sum(ops.values())
    alu = sum(v for k, v in ops.items() if k in ('SHF', 'LOP3', 'IADD3', 'VIADD', 'PRMT', 'ISETP', 'SEL', 'IADD', 'LEA', 'VIADDMNMX', 'IABS', 'MOV'))
    fma = sum(v for k, v in ops.items() if k.startswith('IMAD'))
    print(f"{name[:70]}\n  V={V} total={tot} per-cand={tot/V:.0f}  alu-pipe={alu/V:.0f} fma-pipe={fma/V:.0f}")
    print("   ", ", ".join(f"{k}:{v/V:.0f}" for k, v in ops.most_common(16)))
