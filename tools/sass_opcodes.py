#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libqlinear_b200.so -> profiles/sass_opcodes.txt.

    python tools/sass_opcodes.py [path/to/lib.so] [out.txt]

Evidence that the kernels are Blackwell-native (B200_PROFILING.md "What proves a Blackwell-native
kernel"): tcgen05.mma shows up as UTC*MMA (UTCIMMA = kind::i8, UTCQMMA = kind::f8f6f4, `.2CTA` =
cta_group::2), tcgen05.ld/st as LDTM/STTM, TMA as UTMALDG / UTMASTG / UBLKCP, tcgen05.commit as
UTCBAR; HMMA/IMMA would be the legacy mma.sync path.  Run by __graft_entry__.build() after every
compile, so the committed file always describes the committed sources.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nano-vllm-hw3_b200", "nanovllm", "_lib", "libqlinear_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass_opcodes.txt")
CUOBJDUMP = os.environ.get("CUOBJDUMP", "/usr/local/cuda/bin/cuobjdump")
CUFILT = os.environ.get("CUFILT", "/usr/local/cuda/bin/cu++filt")

FAMILIES = [  # (column, regex on the opcode with its modifiers)
    ("UTCIMMA", r"^UTCIMMA"), ("UTCQMMA", r"^UTCQMMA"), ("UTC*MMA.2CTA", r"^UTC\w*MMA.*\.2CTA"),
    ("UTCBAR", r"^UTCBAR"), ("LDTM", r"^LDTM"), ("STTM", r"^STTM"),
    ("UTMALDG", r"^UTMALDG"), ("UTMASTG", r"^UTMASTG"), ("UTMAPF", r"^UTMAPF|^UTMACCTL"),
    ("SYNCS", r"^SYNCS"), ("legacy HMMA/IMMA", r"^(HMMA|IMMA|QMMA)"), ("LDG", r"^LDG"), ("STG", r"^STG"),
    ("ST peer/.SYS", r"^ST.*\.SYS|^STG.*\.SYS"), ("REDG/ATOMG", r"^(REDG|ATOMG|RED|ATOM)\b"),
]


def demangle(names):
    try:
        r = subprocess.run([CUFILT], input="\n".join(names), capture_output=True, text=True, check=True)
        return r.stdout.splitlines()
    except (OSError, subprocess.CalledProcessError):
        return names


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\((int|bool|unsigned int)\)", "", name)      # template-argument casts
    name = re.sub(r"\(.*$", "", name)
    return name.replace("ql::", "")


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    out = sys.argv[2] if len(sys.argv) > 2 else OUT
    r = subprocess.run([CUOBJDUMP, "-sass", lib], capture_output=True, text=True)
    if r.returncode != 0:
        sys.exit(f"cuobjdump failed: {r.stderr[:400]}")
    kernels, cur = collections.OrderedDict(), None
    arch = set()
    for line in r.stdout.splitlines():
        m = re.match(r"\s*arch = (\w+)", line)
        if m:
            arch.add(m.group(1))
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_x]+)*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    names = list(kernels)
    pretty = [short(n) for n in demangle(names)]
    rows = []
    for n, pn in zip(names, pretty):
        c = kernels[n]
        total = sum(c.values())
        cols = [sum(v for op, v in c.items() if re.search(rx, op)) for _, rx in FAMILIES]
        rows.append((pn, total, cols))
    rows.sort(key=lambda t: t[0])
    with open(out, "w") as f:
        f.write(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}   (arch: {', '.join(sorted(arch))})\n")
        f.write("# per kernel: SASS instruction count, then occurrences of the opcode families below\n")
        f.write("# columns: total | " + " | ".join(n for n, _ in FAMILIES) + "\n")
        for pn, total, cols in rows:
            f.write(f"{pn}\n    {total:6d} | " + " | ".join(f"{v:d}" for v in cols) + "\n")
        tot = [sum(r[2][i] for r in rows) for i in range(len(FAMILIES))]
        f.write(f"TOTAL ({len(rows)} kernels)\n    {sum(r[1] for r in rows):6d} | " + " | ".join(str(v) for v in tot) + "\n")
    print(f"wrote {out}: {len(rows)} kernels")


if __name__ == "__main__":
    main()
