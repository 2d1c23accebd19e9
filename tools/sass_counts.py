#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove the Blackwell paths (B200_PROFILING.md): UTCHMMA (tcgen05.mma),
LDTM / STTM (tcgen05.ld / st: TMEM), UTMALDG (TMA tensor-map load), UBLKCP (1-D bulk async copy), UTCBAR
(tcgen05.commit).  Writes profiles/sass_counts.txt; tests/test_abi.py asserts on the same parser.

    python tools/sass_counts.py [path/to/libb200tts.so]
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "nano-qwen3tts-vllm_b200", "lib", "libb200tts.so")
MNEMONICS = ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "UTCBAR", "HMMA", "SYNCS", "MUFU.EX2")


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.splitlines()
        return out if len(out) == len(names) else names
    except OSError:
        return names


def counts(so=SO):
    out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError(out.stderr)
    res, cur = {}, None
    for line in out.stdout.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            res[cur] = {k: 0 for k in MNEMONICS}
            continue
        if cur is None:
            continue
        for k in MNEMONICS:
            if re.search(r"\b" + re.escape(k) + r"\b", line) or (k + ".") in line:
                res[cur][k] += 1
    names = list(res)
    pretty = demangle(names)
    def short(p):
        p = re.sub(r"\((int|bool|unsigned int)\)", "", p)          # template value casts: <(int)2, (bool)1> -> <2, 1>
        return re.sub(r"\(.*", "", p).replace("void ", "")
    return {short(p): res[n] for n, p in zip(names, pretty)}


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else SO
    c = counts(so)
    lines = [f"# cuobjdump -sass {os.path.relpath(so, ROOT)} : per-kernel mnemonic counts (tools/sass_counts.py)",
             "# " + "  ".join(f"{k:>8}" for k in MNEMONICS) + "  kernel"]
    for name in sorted(c):
        lines.append("  " + "  ".join(f"{c[name][k]:>8}" for k in MNEMONICS) + "  " + name)
    text = "\n".join(lines) + "\n"
    with open(os.path.join(ROOT, "profiles", "sass_counts.txt"), "w") as f:
        f.write(text)
    print(text)


if __name__ == "__main__":
    main()
