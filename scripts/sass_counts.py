"""Per-kernel counts of the SASS mnemonics that prove the sm_100a paths (tcgen05 MMA, TMA, TMEM loads,
packed fp32) are what libavtubes.so contains.  Runs without a GPU:

    python scripts/sass_counts.py > profiles/r2_sass_counts.txt
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "audio_visual_tubes_b200", "libavtubes.so")
MNEMONICS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "LDTM", "SYNCS", "FFMA2", "FMUL2", "FADD2", "MUFU.EX2",
             "MUFU.RCP", "REDG", "RED.", "DFMA", "IDP.4A", "ELECT", "UCGABAR", "STG", "LDG"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    filt = subprocess.run(["c++filt"], input=sass, capture_output=True, text=True).stdout or sass
    funcs, cur = {}, None
    for line in filt.splitlines():
        m = re.search(r"Function : (.+)$", line)
        if m:
            cur = m.group(1).strip()
            funcs[cur] = []
            continue
        if cur is not None and re.search(r"/\*[0-9a-f]{4}\*/", line):
            funcs[cur].append(line)
    arch = re.findall(r"arch = (sm_\w+)", sass)
    print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}: arch {sorted(set(arch))}, {len(funcs)} kernels")
    print("# columns: instructions, then counts of " + ", ".join(MNEMONICS))
    for name, lines in sorted(funcs.items(), key=lambda kv: -len(kv[1])):
        short = name.replace("(anonymous namespace)::", "").replace("avt::", "").replace("void ", "")
        short = re.sub(r"\(.*$", "", short)
        counts = [sum(1 for l in lines if mn in l) for mn in MNEMONICS]
        cells = " ".join(f"{mn}={c}" for mn, c in zip(MNEMONICS, counts) if c)
        print(f"{short:<58s} {len(lines):6d}  {cells}")
    tot = {mn: sum(1 for ls in funcs.values() for l in ls if mn in l) for mn in MNEMONICS}
    print("# library totals: " + " ".join(f"{k}={v}" for k, v in tot.items() if v))


if __name__ == "__main__":
    sys.exit(main())
