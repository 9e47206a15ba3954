"""Per-kernel census of Blackwell-native SASS in libb200llm.so (cuobjdump -sass): tcgen05 MMA (UTC*MMA, .2CTA forms),
TMEM loads / stores (LDTM / STTM), TMA (UTMALDG / UTMASTG / UBLKCP), legacy tensor path (HMMA — expected: none),
128-bit global accesses and warp shuffles. Writes a markdown table (profiles/r2_sass_census.md)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "building_llm_from_scratch_b200", "libb200llm.so")
PATTERNS = [("UTCHMMA", r"\bUTC[A-Z]*MMA"), (".2CTA MMA", r"\bUTC[A-Z]*MMA[.\w]*\.2CTA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"),
            ("UTMALDG", r"\bUTMALDG"), ("UTMASTG", r"\bUTMASTG"), ("UBLKCP", r"\bUBLKCP"), ("UTCBAR", r"\bUTCBAR"),
            ("HMMA (legacy)", r"\bHMMA"), ("LDG.128", r"\bLDG\.E[.\w]*\.128"), ("STG.128", r"\bSTG\.E[.\w]*\.128"),
            ("RED.128", r"\bRED(G)?\.E[.\w]*\.128|\bREDG"), ("SHFL", r"\bSHFL"), ("MUFU.EX2", r"\bMUFU\.EX2"), ("FFMA2", r"\bFFMA2")]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(anonymous namespace\)::", "", cur)
            cur = re.sub(r"\(.*", "", cur).replace("void ", "")
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for name, pat in PATTERNS:
            if re.search(pat, line):
                counts[cur][name] += 1
    cols = [n for n, _ in PATTERNS]
    lines = ["# SASS census of libb200llm.so (sm_100a) — `python tools/sass_census.py`", "",
             "Counts of instruction mnemonics per kernel (`cuobjdump -sass`). `UTC*MMA` = tcgen05.mma, `LDTM`/`STTM` = tcgen05.ld/st,",
             "`UTMALDG`/`UTMASTG` = TMA tile load / store, `UBLKCP` = bulk copy, `HMMA` = legacy mma.sync (none expected).", "",
             "| kernel | " + " | ".join(cols) + " |", "|---|" + "---|" * len(cols)]
    for k, c in counts.items():
        if sum(c.values()) == 0:
            continue
        lines.append(f"| `{k}` | " + " | ".join(str(c[n]) if c[n] else "" for n in cols) + " |")
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    lines.append("| **total** | " + " | ".join(str(tot[n]) for n in cols) + " |")
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r2_sass_census.md")
    open(path, "w").write("\n".join(lines) + "\n")
    print(f"{len(counts)} kernels, totals: " + ", ".join(f"{n} {tot[n]}" for n in cols))


if __name__ == "__main__":
    main()
