#!/usr/bin/env python
"""Write one SASS listing per kernel of the shipped library under profiles/sass/ plus a summary
of the Blackwell-specific mnemonics each one contains (cuobjdump -sass; needs no GPU)."""
import os, re, subprocess, sys, collections

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "p2pfl_b200", "lib", "libp2pfl_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
# opcode prefixes worth counting (matched at the start of the mnemonic)
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKRED", "UTMAPF",
        "UTMACCTL", "SYNCS", "UCGABAR", "FENCE", "LDGMC", "REDG", "ATOMG", "ATOMS", "LDGSTS", "HMMA", "MEMBAR", "MAPA", "STAS", "LDG", "STG",
        "LDS", "STS", "MUFU", "NANOSLEEP"]

def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    parts = re.split(r"\n\s*Function : ", txt)
    names = subprocess.run(["cu++filt"], input="\n".join(p.split("\n", 1)[0].strip() for p in parts[1:]), capture_output=True, text=True).stdout.split("\n")
    rows = []
    for part, dem in zip(parts[1:], names):
        mangled, body = part.split("\n", 1)
        short = re.sub(r"\(.*", "", dem).split("::")[-1].strip() or mangled.strip()
        short = re.sub(r"[^A-Za-z0-9_]", "_", short)[:80]
        lines = [l for l in body.split("\n") if "/*" in l]
        ops = collections.Counter()
        n_instr = 0
        for l in lines:
            m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if not m:
                continue
            n_instr += 1
            op = m.group(1)
            for k in KEYS:
                if op == k or op.startswith(k + ".") or op.startswith(k + "_"):
                    ops[k] += 1
                    break
            if ".STRONG.SYS" in op:
                ops["*.STRONG.SYS (system-scope accesses: peer / multicast memory, flags)"] += 1
        path = os.path.join(OUT, short + ".sass")
        with open(path, "w") as fh:
            fh.write(f"// {dem}\n// cuobjdump -sass of p2pfl_b200/lib/libp2pfl_b200.so (sm_100a), one kernel\n")
            # instruction text only: the hex encodings (half of the file) are dropped
            for l in body.split("\n"):
                if re.match(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", l):
                    continue
                fh.write(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) + "\n")
        rows.append((short, n_instr, ops))
    with open(os.path.join(OUT, "SUMMARY.md"), "w") as fh:
        fh.write("# SASS of every kernel in the shipped library\n\nOne listing per kernel in this directory (`tools/dump_sass.py`, `cuobjdump -sass`, "
                 "sm_100a).  Counts of the mnemonics that show what hardware path a kernel uses: `UTCHMMA` = tcgen05.mma, `LDTM`/`STTM` = "
                 "tcgen05.ld/st (TMEM), `UTMALDG`/`UTMASTG` = TMA tensor load/store, `UBLKCP`/`UBLKRED` = cp.async.bulk / cp.reduce.async.bulk, "
                 "`SYNCS` = mbarrier, `UCGABAR` = cluster barrier (arrive / wait), `FENCE` = proxy fences (generic <-> async), `LDGMC` = multimem.ld_reduce (NVLS), `MAPA` = distributed-shared-memory address mapping; `HMMA` (legacy mma.sync) must be absent.\n\n")
        fh.write("| kernel | instructions | Blackwell / data-plane mnemonics |\n|---|---:|---|\n")
        for short, n, ops in sorted(rows):
            fh.write(f"| `{short}` | {n} | " + ", ".join(f"{k} {v}" for k, v in sorted(ops.items()) if v) + " |\n")
    print(f"{len(rows)} kernels -> {OUT}")

if __name__ == "__main__":
    main()
