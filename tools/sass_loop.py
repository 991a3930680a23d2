"""Instruction mix of the innermost loops that contain a given mnemonic (default HMMA) in a cuobjdump -sass listing.

    cuobjdump -sass -fun <mangled> file.o > k.sass && python tools/sass_loop.py k.sass [MNEMONIC] [--dump]
"""
import collections
import re
import sys


def main():
    path = sys.argv[1]
    key = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else "HMMA"
    ins = []
    for l in open(path):
        m = re.search(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr = {a: i for i, (a, _) in enumerate(ins)}
    loops = []
    for i, (a, t) in enumerate(ins):
        m = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in addr:
            loops.append((addr[int(m.group(1), 16)], i))
    for s, e in sorted(loops, key=lambda se: se[1] - se[0]):
        body = ins[s:e + 1]
        if not any(key in t for _, t in body):
            continue
        c = collections.Counter()
        for _, t in body:
            t = re.sub(r"^@!?U?P\d+\s+", "", t)
            op = t.split()[0]
            c[op if op.startswith("IMAD.MOV") else op.split(".")[0]] += 1
        print(f"loop {ins[s][0]:#x}..{ins[e][0]:#x}: {len(body)} instructions")
        print("  " + ", ".join(f"{k} {v}" for k, v in c.most_common()))
        if "--dump" in sys.argv:
            for a, t in body:
                print(f"  {a:#06x} {t}")
        break


if __name__ == "__main__":
    main()
