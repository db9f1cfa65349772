"""Top SASS lines by warp-stall samples from an ncu report: python scripts/ncu_hot.py report.ncu-rep [kernel-regex] [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"]
if len(sys.argv) > 2 and sys.argv[2] != "-":
    cmd += ["-k", "regex:" + sys.argv[2]]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name",')
for blk in blocks[1:]:
    lines = blk.split("\n")
    print("== kernel:", lines[0][:140])
    rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
    hdr = rows[0]
    si = hdr.index("# Samples"); src = hdr.index("Source")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    body = [r for r in rows[1:] if len(r) == len(hdr)]
    total = sum(int(r[si] or 0) for r in body)
    print("total samples", total, "instructions", len(body))
    idx = sorted(range(len(body)), key=lambda i: -int(body[i][si] or 0))[:n]
    for i in sorted(idx):
        r = body[i]
        st = sorted(((int(r[c] or 0), hdr[c][6:]) for c in stall_cols), reverse=True)[:2]
        print("%5d %5.1f%%  #%4d  %-60s %s" % (int(r[si]), 100.0 * int(r[si]) / max(total, 1), i, r[src].strip()[:60],
                                              " ".join("%s=%d" % (b, a) for a, b in st if a)))
