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
key = (lambda x: -x[0]) if len(sys.argv) > 4 and sys.argv[4] == "inst" else (lambda x: -x[1])
for a in sorted(agg, key=key)[:n]:
    print('%5.1f%% inst %5.1f%% stall  %s:%s  %s' % (100 * a[0] / ti, 100 * a[1] / ts, a[2], a[3], a[4]))
# optional 5th arg: comma-separated line boundaries -> per-range totals for the kernel's main file
if len(sys.argv) > 5:
    bounds = [int(x) for x in sys.argv[5].split(',')]
    files = {}
    for a in agg:
        files.setdefault(a[2], []).append(a)
    for fn, rows in files.items():
        tot = {}
        for ie, ws, _, ln, _ in rows:
            try: ln = int(ln)
            except ValueError: continue
            lo = max([b for b in bounds if b <= ln], default=0)
            t = tot.setdefault(lo, [0, 0]); t[0] += ie; t[1] += ws
        print('file', fn)
        for lo in sorted(tot):
            print('  from line %4d: %5.1f%% inst %5.1f%% stall' % (lo, 100 * tot[lo][0] / ti, 100 * tot[lo][1] / ts))
