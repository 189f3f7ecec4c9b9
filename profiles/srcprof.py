"""Top source lines of a kernel from an .ncu-rep (needs -lineinfo + --import-source on).
usage: python profiles/srcprof.py <report.ncu-rep> <kernel regex> [N]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass',
                      '--kernel-name', 'regex:' + kern], capture_output=True, text=True).stdout
agg, fname, hdr = [], '', None
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == 'File Name': fname = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < len(hdr) or not r[0].strip(): continue
    try:
        ie = int(r[hdr.index('Instructions Executed')] or 0); ws = int(r[hdr.index('Warp Stall Sampling (All Samples)')] or 0)
    except ValueError:
        continue
    agg.append((ie, ws, fname, r[0], r[1].strip()[:100]))
ti = sum(a[0] for a in agg) or 1; ts = sum(a[1] for a in agg) or 1
print(f'kernel {kern}: {ti} warp-instructions, {ts} stall samples')
for a in sorted(agg, key=lambda x: -x[1])[:n]:
    print('%5.1f%% inst %5.1f%% stall  %s:%s  %s' % (100 * a[0] / ti, 100 * a[1] / ts, a[2], a[3], a[4]))
