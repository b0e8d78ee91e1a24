"""Opcode histogram of one kernel's SASS, whole function and its hottest loop (largest backward-branch body).

usage: cuobjdump -sass obj.o > x.sass; python scripts/diag/sass_hist.py x.sass <mangled-name-substring>
"""
import collections
import re
import sys

path, needle = sys.argv[1], sys.argv[2]
lines = open(path).read().split("\n")
start = next(i for i, l in enumerate(lines) if "Function :" in l and needle in l)
end = next((i for i in range(start + 1, len(lines)) if "Function :" in lines[i]), len(lines))
ins = []
for l in lines[start:end]:
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
print("total instructions", len(ins))
loops = []
for addr, text in ins:
    m = re.search(r"\bBRA\S*\s+(?:\S+,\s*)?`?\(?\.?L?_?x?_?(\w+)\)?|BRA\S*\s+.*0x([0-9a-f]+)", text)
    if "BRA" in text:
        t = re.search(r"0x([0-9a-f]+)", text)
        if t:
            tgt = int(t.group(1), 16)
            if tgt < addr:
                loops.append((addr - tgt, tgt, addr))
loops.sort(reverse=True)
for size, tgt, addr in loops[:3]:
    body = [t for a, t in ins if tgt <= a <= addr]
    hist = collections.Counter(re.sub(r"^@!?\w+\s+", "", t).split()[0].split(".")[0] for t in body)
    print(f"loop {tgt:#x}..{addr:#x}: {len(body)} instructions")
    print("  ", ", ".join(f"{k} {v}" for k, v in hist.most_common(24)))
