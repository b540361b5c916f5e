"""Full SASS listings of the shipped library, one file per kernel family, plus the mnemonic evidence table.

    python scripts/dump_sass.py            # needs only cuobjdump (no GPU)

`profiles/sass/<source>.sass` holds the disassembly of every kernel compiled from csrc/<source>.cu (function by function, as
cuobjdump prints it for sm_100a); `profiles/sass/summary.md` counts the mnemonics that identify the Blackwell paths
(B200_PROFILING.md): UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA, LDGMC/STGMC... = multimem,
.SYS-scoped LDG/STG with STRONG = the release/acquire flag traffic of the fused collectives.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "neuroimagedisttraining_b200", "build", "obj")
OUT = os.path.join(ROOT, "profiles", "sass")
PAT = collections.OrderedDict([
    ("UTC*MMA (tcgen05.mma)", r"\bUTC\w*MMA"), ("LDTM (tcgen05.ld)", r"\bLDTM"), ("STTM (tcgen05.st)", r"\bSTTM"),
    ("UTMALDG (TMA load)", r"\bUTMALDG"), ("UTMASTG (TMA store)", r"\bUTMASTG"), ("UBLKCP (bulk copy)", r"\bUBLKCP"),
    ("UTCBAR (tcgen05.commit)", r"\bUTCBAR"), ("SYNCS (mbarrier)", r"\bSYNCS"),
    ("multimem ld_reduce (LDGMC/..RED)", r"\bLDGMC|\bLDG\S*\.MC"), ("multimem st / red (STG.MC / REDG.MC)", r"\bSTGMC|\bSTG\S*\.MC|\bREDG\S*\.MC|\bREDGMC"),
    ("ld.acquire.sys (LDG..STRONG.SYS)", r"\bLDG\S*STRONG\.SYS"), ("st.release.sys (STG..STRONG.SYS)", r"\bSTG\S*STRONG\.SYS"),
    ("HMMA (legacy mma.sync)", r"\bHMMA"), ("FHFMA.BF16 (fma.rn.f32.bf16, mixed precision)", r"\bFHFMA"),
])


def main():
    os.makedirs(OUT, exist_ok=True)
    rows = []
    for obj in sorted(f for f in os.listdir(OBJ) if f.endswith(".o")):
        src = obj[:-2]
        r = subprocess.run(["cuobjdump", "-sass", os.path.join(OBJ, obj)], capture_output=True, text=True)
        if r.returncode != 0:
            print("cuobjdump failed for", obj, r.stderr[-200:], file=sys.stderr)
            continue
        # drop the hex encodings (trailing comment + the encoding-only continuation lines): a third of the size, same instructions
        text = re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/", "", r.stdout)
        text = "\n".join(l for l in text.split("\n") if l.strip())
        with open(os.path.join(OUT, src + ".sass"), "w") as f:
            f.write(text)
        funcs = re.split(r"\n\s*Function : ", text)[1:]
        for fn in funcs:
            name = fn.split("\n", 1)[0].strip()
            dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
            short = re.sub(r"\(.*", "", dem)[-70:]
            n_instr = len(re.findall(r"^\s+/\*[0-9a-f]{4}\*/", fn, flags=re.M))
            counts = [len(re.findall(p, fn)) for p in PAT.values()]
            rows.append((src, short, n_instr, counts))
    with open(os.path.join(OUT, "summary.md"), "w") as f:
        f.write("# SASS evidence per kernel (cuobjdump -sass of the objects linked into lib/libnidt_b200.so, sm_100a)\n\n")
        f.write("Full listings: `profiles/sass/<source>.sass`.  Columns count instruction mnemonics per kernel.\n\n")
        f.write("| source | kernel | instr | " + " | ".join(PAT) + " |\n|---|---|---|" + "---|" * len(PAT) + "\n")
        for src, short, n, counts in rows:
            f.write(f"| {src}.cu | `{short}` | {n} | " + " | ".join(str(c) if c else "" for c in counts) + " |\n")
    print(f"{len(rows)} kernels -> {OUT}")


if __name__ == "__main__":
    main()
