"""Instruction mix of every loop (backward branch) in a cuobjdump -sass listing.

    cuobjdump -sass -fun <mangled kernel> lib.so | python tools/sass_loops.py
"""
import re
import sys

ins = []
for l in sys.stdin:
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
addr = {a: i for i, (a, _) in enumerate(ins)}
loops = []
for i, (a, t) in enumerate(ins):
    if 'BRA' in t:
        m = re.search(r'0x([0-9a-f]+)', t)
        if m:
            tg = int(m.group(1), 16)
            if tg < a and tg in addr:
                loops.append((addr[tg], i))
print(len(ins), "instructions;", len(loops), "loops")
for s, e in loops:
    body = [t for _, t in ins[s:e+1]]
    c = lambda p: sum(1 for t in body if re.search(p, t))
    print(f"{ins[s][0]:#x}-{ins[e][0]:#x} n={e-s+1} D={c(r'D(ADD|MUL|FMA)')} F2F={c('F2F')} "
          f"MUFU={c('MUFU')} FP2={c(r'F(FMA|MUL|ADD)2')} FP={c(r'F(FMA|MUL|ADD) ')} LDS={c('LDS')} "
          f"MOV={c(r'MOV')} BRA={c('BRA|BSSY|BSYNC')}")
