#!/usr/bin/env python3
"""Prints the innermost hot loop of a kernel from `cuobjdump -sass -fun NAME lib.so`:
the backward-branch loop holding the most instructions matching PATTERN (default: packed
FP32), with a per-mnemonic count.  Usage: sass_hot_loop.py file.sass [regex]"""
import collections
import re
import signal
import sys

signal.signal(signal.SIGPIPE, signal.SIG_DFL)


def main():
    path = sys.argv[1]
    pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else r"FFMA2|FADD2|FMUL2")
    ins = []
    for line in open(path):
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    addr_index = {a: i for i, (a, _) in enumerate(ins)}
    best = None
    for i, (a, text) in enumerate(ins):
        m = re.search(r"\bBRA\b.*?\b0x([0-9a-f]+)", text)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt >= a or tgt not in addr_index:
            continue
        body = ins[addr_index[tgt]:i + 1]
        hits = sum(1 for _, t in body if pat.search(t))
        # innermost = densest: prefer the smaller loop on ties
        key = (hits / max(len(body), 1), hits)
        if hits and (best is None or key > best[0]):
            best = (key, body)
    if best is None:
        print("no loop with", pat.pattern)
        return 1
    body = best[1]
    count = collections.Counter()
    for _, t in body:
        t = re.sub(r"^@!?U?P\d+\s+", "", t)
        count[t.split()[0].split(".")[0]] += 1
    print("# loop 0x%04x..0x%04x: %d instructions" % (body[0][0], body[-1][0], len(body)))
    print("# " + ", ".join("%s %d" % kv for kv in count.most_common()))
    for a, t in body:
        print("/*%04x*/ %s" % (a, t))
    return 0


if __name__ == "__main__":
    sys.exit(main())
