"""SASS mnemonic counts per kernel of the shipped library: python tools/sass_counts.py > profiles/<round>_sass_counts.md
UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA load / store, LDTM / STTM = tcgen05.ld / st, HMMA = legacy mma.sync,
REDG / ATOMG = global reductions / atomics."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "stable_diffusion_training_b200", "libsdb200.so")
OPS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "HMMA", "MUFU", "REDG", "ATOMG"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return dict(zip(names, out))


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts, lines, cur = collections.OrderedDict(), {}, None
    for ln in sass.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            lines[cur] = 0
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
        if m:
            lines[cur] += 1
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    counts[cur][o] += 1
    dm = demangle(list(counts))
    short = lambda n: re.sub(r"\(.*", "", dm[n].replace("(anonymous namespace)::", ""))
    print("# SASS mnemonic counts per kernel of the shipped libsdb200.so\n")
    print("`python tools/sass_counts.py` (`cuobjdump -sass stable_diffusion_training_b200/libsdb200.so`, counted per "
          "`Function :` section). UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA load / store, LDTM / STTM = tcgen05.ld / "
          "st, HMMA = the legacy mma.sync path, REDG / ATOMG = global reductions / atomics.\n")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print("Library totals: " + ", ".join(f"{o} {tot[o]}" for o in OPS) + f"; {len(counts)} kernels.\n")
    print("| kernel | " + " | ".join(OPS) + " | SASS instructions |")
    print("|---|" + "---|" * (len(OPS) + 1))
    order = sorted(counts, key=lambda n: (-(counts[n]["UTCHMMA"] > 0), -lines[n]))
    for n in order:
        print(f"| `{short(n)}` | " + " | ".join(str(counts[n][o]) for o in OPS) + f" | {lines[n]} |")


if __name__ == "__main__":
    sys.exit(main())
