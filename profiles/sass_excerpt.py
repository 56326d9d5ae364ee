"""SASS evidence for the kernels of libkafka_b200.so (run where the library is built; needs cuobjdump):

    python profiles/sass_excerpt.py > profiles/r02_sass_excerpt.txt

* mnemonic counts over the whole library: bulk-copy (TMA) and cp.async instructions, fp64 arithmetic,
  and the absence of any tensor-core instruction (the path is fp64 vector arithmetic);
* the hot loop of gp_linearize_kernel<10,2> (one trip = one training point for two pixels);
* the band loop of solve_bands_kernel<10,1> (one trip = one band of one pixel's normal equations).
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kafka_inferenceengine_b200", "csrc", "libkafka_b200.so")


def functions(text):
    out, name, rows = {}, None, []
    for line in text.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                out[name] = rows
            name, rows = m.group(1), []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and name:
            rows.append((int(m.group(1), 16), m.group(2).strip()))
    if name:
        out[name] = rows
    return out


def mnemonic(ins):
    parts = ins.split()
    op = parts[1] if parts[0].startswith("@") else parts[0]
    return op.split(".")[0]


def innermost_loop(rows, want):
    """The backward branch whose body holds the most `want` instructions."""
    best = None
    for addr, ins in rows:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?0x([0-9a-f]+)", ins)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt >= addr:
            continue
        body = [(a, i) for a, i in rows if tgt <= a <= addr]
        n = sum(1 for _, i in body if mnemonic(i) == want)
        inner = False                        # another backward branch inside: not the innermost loop
        for a, i in body:
            mm = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d+,\s*)?0x([0-9a-f]+)", i)
            if mm and a != addr and int(mm.group(1), 16) < a:
                inner = True
        if n and not inner and (best is None or n > best[0]):
            best = (n, body)
    return best[1] if best else []


def main():
    text = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    fns = functions(text)
    total = collections.Counter()
    for rows in fns.values():
        total.update(mnemonic(i) for _, i in rows)
    print("libkafka_b200.so: %d kernels, sm_100a SASS" % len(fns))
    print("mnemonic counts over the whole library:")
    for op in ("UBLKCP", "LDGSTS", "SYNCS", "DFMA", "DADD", "DMUL", "MUFU", "SHFL", "LDS", "STS", "LDG", "STG", "LDL", "STL"):
        print("  %-8s %7d" % (op, total.get(op, 0)))
    tensor = {k: v for k, v in total.items() if re.match(r"(HMMA|IMMA|DMMA|QMMA|UTC|UTCHMMA|UTCQMMA|BMMA|HGMMA)", k)}
    print("  tensor-core instructions (HMMA/IMMA/DMMA/UTC*MMA/...): %s" % (tensor or "none"))
    for pat, want, title in (("gp_linearize_kernelILi10ELi2", "DFMA", "gp_linearize_kernel<10,2>: loop over training points"),
                             ("solve_bands_kernelILi10ELi1", "DFMA", "solve_bands_kernel<10,1>: loop over bands")):
        name = next(n for n in fns if pat in n)
        rows = fns[name]
        c = collections.Counter(mnemonic(i) for _, i in rows)
        body = innermost_loop(rows, want)
        cb = collections.Counter(mnemonic(i) for _, i in body)
        print("\n%s\n  %s\n  whole kernel: %d instructions, %s" % (title, name, len(rows),
              ", ".join("%s %d" % kv for kv in c.most_common(8))))
        print("  loop body: %d instructions, %s" % (len(body), ", ".join("%s %d" % kv for kv in cb.most_common(12))))
        for a, i in body:
            print("    /*%04x*/ %s" % (a, i))


if __name__ == "__main__":
    main()
