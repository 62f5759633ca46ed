"""SASS evidence for profiles/: for every kernel in libdtg_b200.so, an excerpt of the actual
`cuobjdump -sass` listing around the instructions that matter (tcgen05 = UTC*MMA, tcgen05.ld =
LDTM, TMA = UTMALDG / UTMASTG / UTMAREDG / UBLKCP, remote reductions = REDG / ATOMG with .SYS,
NVLS multimem = *MC*), plus mnemonic counts.  Runs on the CPU box (no GPU needed).

    python tools/sass_evidence.py            -> profiles/sass/<kernel>.sass + profiles/sass/INDEX.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-tensorflow-guide_b200", "ops", "libdtg_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEY = re.compile(r"\b(UTC\w*MMA\S*|UTCBAR\S*|LDTM\S*|STTM\S*|UTMALDG\S*|UTMASTG\S*|UTMAREDG\S*|UBLKCP\S*|"
                 r"UTMAPF\S*|UTCCP\S*|SYNCS\S*|REDG\S*|ATOMG\S*|RED\.\S*|LDG\S*MC\S*|STG\S*MC\S*|"
                 r"\S*MULTIMEM\S*|LDGMC\S*|REDUX\S*|MUFU\S*|HMMA\S*|UCGABAR\S*|CCTL\S*|MEMBAR\S*)")
CTX = 2
PRIORITY = ("UTC", "LDTM", "STTM", "UTMA", "UBLKCP", "LDGMC", "STGMC", "REDG", "ATOMG", "HMMA")


def demangle(names):
    r = subprocess.run(["c++filt"] + names, stdout=subprocess.PIPE, text=True)
    return r.stdout.strip().splitlines()


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], stdout=subprocess.PIPE, text=True).stdout
    os.makedirs(OUT, exist_ok=True)
    blocks = re.split(r"\n\s*Function : ", sass)[1:]
    names = [b.split("\n", 1)[0].strip() for b in blocks]
    pretty = demangle(names)
    index = ["# SASS evidence (cuobjdump -sass of libdtg_b200.so, sm_100a)\n",
             "One file per kernel: the instruction lines that prove the hardware path, with %d lines "
             "of context, taken verbatim from the listing.  Regenerate with `python tools/sass_evidence.py`.\n" % CTX,
             "| kernel | instructions | key mnemonics (count) | file |", "|---|---|---|---|"]
    seen = collections.Counter()
    for name, nice, body in zip(names, pretty, blocks):
        lines = [l for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
        hits = [i for i, l in enumerate(lines) if KEY.search(l)]
        cnt = collections.Counter()
        for i in hits:
            m = KEY.search(lines[i]).group(1)
            cnt[re.sub(r"\..*", "", m) if not m.startswith(("REDG", "ATOMG", "UTC")) else
                ".".join(m.split(".")[:4])] += 1
        short = re.sub(r"\(.*", "", nice).replace("void ", "").replace("dtg::", "")
        short = re.sub(r"[^A-Za-z0-9_<>,]", "_", short)[:80]
        seen[short] += 1
        fn = "%s%s.sass" % (short.replace("<", "_").replace(">", "").replace(",", "_"),
                            "" if seen[short] == 1 else "_%d" % seen[short])
        keep = set()
        # cap the excerpt: first 40 hits of each mnemonic class
        per = collections.Counter()
        for i in hits:
            k = KEY.search(lines[i]).group(1).split(".")[0]
            per[k] += 1
            if per[k] <= 40:
                keep.update(range(max(0, i - CTX), min(len(lines), i + CTX + 1)))
        with open(os.path.join(OUT, fn), "w") as f:
            f.write("// %s\n// %d SASS instructions; excerpt around: %s\n" % (
                nice, len(lines), ", ".join("%s x%d" % kv for kv in cnt.most_common(12))))
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    f.write("        ...\n")
                f.write(lines[i].rstrip() + "\n")
                last = i
        # the hardware-path mnemonics first (tensor core, TMEM, TMA, multimem, remote reductions),
        # then whatever else is frequent
        first = [kv for kv in cnt.most_common() if kv[0].startswith(PRIORITY)]
        rest = [kv for kv in cnt.most_common() if not kv[0].startswith(PRIORITY)]
        shown = first + rest[:max(0, 8 - len(first))]
        index.append("| `%s` | %d | %s | [%s](%s) |" % (
            nice[:110].replace("|", "\\|"), len(lines),
            ", ".join("%s (%d)" % kv for kv in shown) or "-", fn, fn))
    with open(os.path.join(OUT, "INDEX.md"), "w") as f:
        f.write("\n".join(index) + "\n")
    print("wrote %d kernels to %s" % (len(names), OUT))


if __name__ == "__main__":
    sys.exit(main())
