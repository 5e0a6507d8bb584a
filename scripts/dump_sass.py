"""Regenerate profiles/sass/: one full `cuobjdump -sass` listing per kernel of the shipped libb200cnn.so, plus INDEX.md
with, per kernel, the counts of the opcodes that prove what it runs on (B200_PROFILING.md "What proves a Blackwell-native
kernel"): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA, LDGMC = multimem.ld_reduce (in-switch
reduction), STG/LDG ...STRONG.SYS = system-scope peer accesses, HMMA = legacy mma.sync (must be absent).

    python scripts/dump_sass.py            # no GPU needed
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "azure_hc_intel_tf_b200", "_lib", "libb200cnn.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "LDGMC", "SYNCS", "UCGABAR", "FHFMA", "HMMA", "MEMBAR"]


def demangle(names):
    p = subprocess.run(["cu++filt"] + names, capture_output=True, text=True)
    out = p.stdout.strip().splitlines()
    return out if len(out) == len(names) else names


def short(dem):
    s = re.sub(r"b200::\(anonymous namespace\)::|b200::<unnamed>::|void ", "", dem)
    s = re.sub(r"\((bool|int|unsigned int|long)\)", "", s)   # (bool)1 -> 1 in template arguments
    depth, cut = 0, len(s)
    for i, ch in enumerate(s):                       # drop the argument list: first '(' outside template brackets
        if ch == "<":
            depth += 1
        elif ch == ">":
            depth -= 1
        elif ch == "(" and depth == 0:
            cut = i
            break
    s = s[:cut]
    s = re.sub(r"[^A-Za-z0-9_<>,]+", "_", s).strip("_")
    return s.replace("<", "_").replace(">", "").replace(",", "_")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"^\s*Function : (\S+)\s*$", sass, flags=re.M)
    os.makedirs(OUT, exist_ok=True)
    for f in os.listdir(OUT):
        if f.endswith(".sass"):
            os.remove(os.path.join(OUT, f))
    names = parts[1::2]
    bodies = parts[2::2]
    dems = demangle(names)
    rows = []
    for mangled, dem, body in zip(names, dems, bodies):
        ops = collections.Counter()
        n_instr = 0
        sys_scope = 0
        for line in body.splitlines():
            m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Za-z0-9_]+)*)", line)
            if not m:
                continue
            n_instr += 1
            op = m.group(1)
            for k in KEYS:
                if op.startswith(k):
                    ops[k] += 1
            if ".SYS" in m.group(2) and op in ("LDG", "STG", "LDGMC", "ATOMG", "REDG"):
                sys_scope += 1
        # keep address + instruction; drop the 128-bit encodings (they triple the size and nobody reads them)
        body = re.sub(r"[ \t]*/\* 0x[0-9a-f]{16} \*/", "", body)
        body = "\n".join(l.rstrip() for l in body.splitlines() if l.strip())
        fn = short(dem)[:110] + ".sass"
        with open(os.path.join(OUT, fn), "w") as f:
            f.write(f"// {dem}\n// {mangled}\n// from libb200cnn.so (sm_100a), cuobjdump -sass\n{body}")
        rows.append((short(dem), fn, n_instr, ops, sys_scope))
    rows.sort()
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("# SASS listings of every kernel in libb200cnn.so (sm_100a)\n\n"
                "Regenerate with `python scripts/dump_sass.py` after a build. Opcode columns: UTCHMMA = tcgen05.mma, "
                "LDTM = tcgen05.ld, UTMALDG / UTMASTG = TMA load / store (`.IM2COL`, `.2CTA` variants included), "
                "LDGMC = multimem.ld_reduce (NVSwitch in-switch reduction; multimem.st assembles to a STG on the "
                "multicast address), `.SYS` = loads / stores / reductions at system scope (peer memory, flags), "
                "UCGABAR = cluster barrier, FHFMA = mixed-precision bf16 x bf16 + fp32 FMA (the depthwise kernels), HMMA = legacy "
                "mma.sync (none anywhere).\n\n")
        f.write("| kernel | file | instr | " + " | ".join(KEYS) + " | .SYS ld/st |\n|---|---|---:|" + "---:|" * (len(KEYS) + 1) + "\n")
        for name, fn, n, ops, sysn in rows:
            f.write(f"| `{name[:80]}` | [{fn[:40]}…]({fn}) | {n} | " + " | ".join(str(ops.get(k, 0) or "") for k in KEYS) + f" | {sysn or ''} |\n")
    print(f"wrote {len(rows)} listings to {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
