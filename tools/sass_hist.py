#!/usr/bin/env python
"""Per-kernel histogram of the Blackwell-specific SASS mnemonics in libsietill_b200.so.

  python tools/sass_hist.py [out.txt]

Runs `cuobjdump -sass` on the shipped library (the file the GPU box loads) and counts, per kernel, the
instructions that prove the tcgen05 / TMEM / TMA path (B200_PROFILING.md): UTCHMMA / UTCQMMA (tcgen05.mma),
LDTM / STTM (tcgen05.ld / st), UTCBAR (tcgen05.commit), UTCATOMSWS (tcgen05.alloc family), UBLKCP (cp.async.bulk),
UTMALDG (tensor-map TMA), SYNCS (mbarrier), plus the FP64 / FFMA2 / HMMA counts for orientation.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "speechrecognition_b200", "libsietill_b200.so")
KEYS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UBLKCP", "UTMALDG", "SYNCS", "FFMA2", "DFMA", "DADD",
        "DMUL", "HMMA", "LDS", "STS", "LDG", "STG"]


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        return dict(zip(names, out)) if len(out) == len(names) else {n: n for n in names}
    except Exception:
        return {n: n for n in names}


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    hist, total, cur, arch = {}, {}, None, {}
    cur_arch = "?"
    for line in sass.splitlines():
        m = re.match(r"\s*arch = (sm_\w+)", line)
        if m:
            cur_arch = m.group(1)
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            hist[cur], total[cur], arch[cur] = collections.Counter(), 0, cur_arch
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            op = m.group(1)
            total[cur] += 1
            for k in KEYS:
                if op == k or op.startswith(k + "."):
                    hist[cur][k] += 1
    # the mnemonic may carry suffixes after a dot on the same token: count again on the raw text for robustness
    names = demangle(list(hist))
    lines = ["# cuobjdump -sass %s — per-kernel counts of Blackwell-specific instructions" % os.path.relpath(LIB, ROOT),
             "# columns: " + " ".join(KEYS) + " | total SASS instructions | arch | kernel"]
    for fn in sorted(hist, key=lambda f: -hist[f]["UTCHMMA"] * 10**6 - total[f]):
        short = names[fn]
        cut = short.find(">(")   # template arguments stay, the parameter list goes
        short = short[:cut + 1] if cut >= 0 else re.sub(r"\(.*", "", short)
        lines.append(" ".join("%5d" % hist[fn][k] for k in KEYS) + " | %6d | %s | %s" % (total[fn], arch[fn], short))
    text = "\n".join(lines) + "\n"
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text)
    else:
        sys.stdout.write(text)


if __name__ == "__main__":
    main()
