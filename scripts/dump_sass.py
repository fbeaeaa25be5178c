"""Per-kernel SASS evidence for profiles/sass/: instruction histogram + the listing around the tcgen05 / TMA / TMEM
instructions of every kernel in libdedloc_b200.so (full listings: `cuobjdump -sass dedloc_b200/lib/libdedloc_b200.so`).

    python scripts/dump_sass.py            # writes profiles/sass/<kernel>.txt and profiles/sass/INDEX.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dedloc_b200", "lib", "libdedloc_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEY = re.compile(r"LDGMC|MULTIMEM|\.MC|MEMBAR|FENCE|ERRBAR|CCTL|UTCHMMA|UTMALDG|UTMASTG|UTMAPF|LDTM|STTM|UTCBAR|UTCCP|SYNCS|UCGABAR|MUFU|REDG|RED\.|ATOMG|LDG\.E\.128|STG\.E\.128|LD\.E\.128")


def demangle(name: str) -> str:
    try:
        return subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    except OSError:
        return name


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    os.makedirs(OUT, exist_ok=True)
    index = ["# SASS evidence per kernel (sm_100a)", "",
             "| kernel | instructions | UTCHMMA | UTMALDG | UTMASTG | LDTM | STTM | MUFU | file |", "|---|---|---|---|---|---|---|---|---|"]
    funcs = re.split(r"\n\s*Function : ", sass)[1:]
    for blk in funcs:
        mangled, _, body = blk.partition("\n")
        mangled = mangled.strip()
        lines = [re.sub(r"/\* 0x[0-9a-f]+ \*/", "", l).rstrip() for l in body.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,5}\*/", l)]
        ops = []
        for l in lines:
            m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
            if m:
                ops.append(m.group(1))
        hist = collections.Counter(o.split(".")[0] for o in ops)
        pretty = demangle(mangled)
        short = re.sub(r"[^A-Za-z0-9_]+", "_", re.sub(r"\(.*", "", pretty).replace("dedloc::", ""))[:80].strip("_")
        path = os.path.join(OUT, short + ".txt")
        with open(path, "w") as f:
            f.write(f"kernel: {pretty}\nmangled: {mangled}\ninstructions: {len(ops)}\n\nhistogram (top 30):\n")
            for k, v in hist.most_common(30):
                f.write(f"  {k:12s} {v}\n")
            f.write("\nlisting around the tensor-core / TMA / TMEM / async-barrier instructions (first 160 matching lines, +-1 line):\n")
            shown, last = 0, -10
            for i, l in enumerate(lines):
                if KEY.search(l):
                    lo = max(i - 1, last + 1)
                    for j in range(lo, min(i + 2, len(lines))):
                        f.write(lines[j] + "\n")
                    last = i + 1
                    shown += 1
                    if shown >= 160:
                        f.write("  ...\n")
                        break
        index.append(f"| `{pretty[:90]}` | {len(ops)} | {hist['UTCHMMA']} | {hist['UTMALDG']} | {hist['UTMASTG']} | {hist['LDTM']} | "
                     f"{hist['STTM']} | {hist['MUFU']} | `sass/{short}.txt` |")
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("\n".join(index) + "\n")
    print(f"{len(funcs)} kernels -> {OUT}")


if __name__ == "__main__":
    sys.exit(main())
