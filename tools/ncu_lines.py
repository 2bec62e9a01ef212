"""Per-source-line share of stall samples and executed instructions of one kernel in an .ncu-rep
(needs -lineinfo and --import-source on).  usage: ncu_lines.py report.ncu-rep <kernel id> [top]"""
import csv, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid, "--print-source", "cuda,sass"],
                     stdout=subprocess.PIPE, text=True).stdout
rows = list(csv.reader(out.splitlines()))
print(rows[1][1])
hdr = rows[2]
isamp, iinst, ithr = hdr.index('# Samples'), hdr.index('Instructions Executed'), hdr.index('Thread Instructions Executed')
agg = {}
for r in rows[3:]:
    if len(r) < 10 or r[2] != '-':
        continue
    try:
        agg[int(r[0])] = (r[1], int(r[isamp] or 0), int(r[iinst] or 0), int(r[ithr] or 0))
    except ValueError:
        pass
ts, ti = sum(v[1] for v in agg.values()), sum(v[2] for v in agg.values())
print("total samples", ts, "warp instructions", ti)
key = (lambda kv: -kv[1][2]) if len(sys.argv) > 4 else (lambda kv: -kv[1][1])
for ln, (s, sa, i, t) in sorted(agg.items(), key=key)[:top]:
    print(f"{ln:4d} samp {100*sa/ts:5.1f}% inst {100*i/ti:5.1f}% thr/inst {t/max(i,1):5.1f} | {s.strip()[:120]}")
