"""Top source lines of each kernel in an .ncu-rep by stall samples / executed instructions (needs -lineinfo + --import-source on)."""
import csv, io, subprocess, sys
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 14
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
kern, hdr, cur = None, None, []
def flush():
    if not cur:
        return
    i_s, i_e = hdr.index("# Samples"), hdr.index("Instructions Executed")
    def num(r, i):
        try:
            return int(r[i] or 0)
        except (ValueError, IndexError):
            return 0
    tot_s = sum(num(r, i_s) for r in cur) or 1
    tot_e = sum(num(r, i_e) for r in cur) or 1
    print(f"\n## {kern[:110]}\n   total samples {tot_s}, warp instructions {tot_e}")
    for r in sorted(cur, key=lambda r: -num(r, i_s))[:topn]:
        print(f"   {100*num(r, i_s)/tot_s:5.1f}% smp {100*num(r, i_e)/tot_e:5.1f}% ins  L{r[0]:>4}: {r[1].strip()[:120]}")
for r in rows:
    if len(r) >= 2 and r[0] == "Function Name":
        flush(); kern, cur = r[1], []
    elif r and r[0] == "Line No":
        hdr = r
    elif kern and hdr and len(r) > 8 and r[0] not in ("", "File Path"):
        cur.append(r)
flush()
