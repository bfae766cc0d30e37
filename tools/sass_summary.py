"""Per-object SASS evidence for profiles/sass_summary.txt: counts of the Blackwell tensor-core / TMEM / TMA mnemonics
in every object of libb200_hotpath.so (cuobjdump -sass; B200_PROFILING.md lists the mnemonics).

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATTERNS = [("UTC*MMA (tcgen05.mma)", r"\bUTC[A-Z0-9]*MMA"), ("LDTM/STTM (tcgen05.ld/st)", r"\b(LDTM|STTM)"),
            ("UTMALDG/UTMASTG (TMA tensor)", r"\bUTMA(LDG|STG)"), ("UBLKCP (bulk copy)", r"\bUBLKCP"),
            ("SYNCS (mbarrier)", r"\bSYNCS"), ("LDGSTS (cp.async)", r"\bLDGSTS"), ("HMMA/IMMA (mma.sync)", r"\b(HMMA|IMMA)"),
            ("UTCBAR (tcgen05.commit)", r"\bUTCBAR")]


def main():
    objs = sorted(glob.glob(os.path.join(ROOT, "sglang_npu_b200", "lib", "obj", "*.o")))
    print("# cuobjdump -sass per object of sglang_npu_b200/lib/libb200_hotpath.so (sm_100a), instruction counts")
    print("%-24s" % "object" + "".join("%30s" % n for n, _ in PATTERNS))
    for o in objs:
        sass = subprocess.run(["cuobjdump", "-sass", o], capture_output=True, text=True).stdout
        arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
        row = "%-24s" % os.path.basename(o)
        for _, pat in PATTERNS:
            row += "%30d" % len(re.findall(pat, sass))
        print(row + "   " + ",".join(arch))


if __name__ == "__main__":
    main()
