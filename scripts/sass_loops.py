#!/usr/bin/env python
"""Static SASS loop report: for every backward branch in a kernel, the loop body length and opcode mix.
Usage: sass_loops.py <cubin|so> <kernel-name-substring> [min_len]"""
import collections
import re
import subprocess
import sys


def main(path, needle, min_len=20):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", out)
    for f in funcs[1:]:
        name = f.split("\n", 1)[0]
        if needle not in name:
            continue
        ins = []
        for line in f.split("\n"):
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
            if m:
                ins.append((int(m.group(1), 16), m.group(2).strip()))
        addr_index = {a: i for i, (a, _) in enumerate(ins)}
        print("==", name, "instructions:", len(ins))
        for i, (a, text) in enumerate(ins):
            m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:!?U?P\d+,\s*)?`?\(?\.?(?:L_x_\d+|0x([0-9a-f]+))", text)
            if not m or not m.group(1):
                continue
            tgt = int(m.group(1), 16)
            if tgt <= a and tgt in addr_index:
                body = ins[addr_index[tgt]:i + 1]
                if len(body) < min_len:
                    continue
                c = collections.Counter()
                for _, t in body:
                    sp = t.split()
                    op = sp[1] if sp[0].startswith("@") else sp[0]
                    c[op.split(".")[0] if not op.startswith(("LDS", "LDL", "STL", "LDG", "STG", "DMMA")) else op] += 1
                fp64 = sum(v for k, v in c.items() if k in ("DMUL", "DADD", "DFMA", "DSETP"))
                print("  loop @%05x len %4d  fp64 %3d  dmma %2d  lds %2d  ldl/stl %2d | %s" % (
                    tgt, len(body), fp64, sum(v for k, v in c.items() if k.startswith("DMMA")),
                    sum(v for k, v in c.items() if k.startswith("LDS")), sum(v for k, v in c.items() if k.startswith(("LDL", "STL"))),
                    " ".join("%s:%d" % kv for kv in c.most_common(9))))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 20)
