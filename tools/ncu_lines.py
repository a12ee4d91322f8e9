"""Per-source-line totals of an ncu capture: joins `ncu --page source --csv` (SASS rows) with `nvdisasm --print-line-info-inline`.
    python tools/ncu_lines.py <rep.ncu-rep> <kernel regex> <object.o> <mangled-name substring> <source file basename> [top]
Prints, per line of the source file: warp instructions executed, share, stall samples, share, shared-memory wavefronts."""
import csv, os, re, subprocess, sys, tempfile
from collections import defaultdict

rep, kre, obj, mangled, srcname = sys.argv[1:6]
top = int(sys.argv[6]) if len(sys.argv) > 6 else 40
tmp = tempfile.mkdtemp()
subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
dis = subprocess.run(['nvdisasm', '--print-line-info-inline', os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
lines_of = {}
infn, cur = False, None
pend = []
for l in dis.splitlines():
    if l.startswith('.text.'):
        infn = mangled in l
        continue
    if not infn:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        pend.append((os.path.basename(m.group(1)), int(m.group(2))))
        continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*);', l)
    if m:
        if pend:
            mine = [p for p in pend if p[0] == srcname]
            cur = mine[0][1] if mine else cur
            pend = []
        lines_of[int(m.group(1), 16)] = cur
csvtxt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(csvtxt.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]
ci = {n: hdr.index(n) for n in ('Address', 'Source', '# Samples', 'Instructions Executed', 'L1 Wavefronts Shared', 'Thread Instructions Executed')}
body = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or not r[0].startswith('0x'):
        if r and r[0] == 'Kernel Name':
            break                      # first matching launch only
        continue
    body.append(r)
base = int(body[0][ci['Address']], 16)
ins, smp, wav, thr = defaultdict(float), defaultdict(float), defaultdict(float), defaultdict(float)
for r in body:
    off = int(r[ci['Address']], 16) - base
    ln = lines_of.get(off)
    ins[ln] += float(r[ci['Instructions Executed']] or 0)
    smp[ln] += float(r[ci['# Samples']] or 0)
    wav[ln] += float(r[ci['L1 Wavefronts Shared']] or 0)
    thr[ln] += float(r[ci['Thread Instructions Executed']] or 0)
TI, TS = sum(ins.values()), sum(smp.values())
src = open(os.path.join(os.path.dirname(os.path.abspath(obj)), srcname)).read().splitlines()
print('kernel %s: %.3g warp instructions, %.0f samples, %.3g shared wavefronts, avg active threads %.1f' % (kre, TI, TS, sum(wav.values()), sum(thr.values()) / max(TI, 1)))
for ln in sorted(ins, key=lambda k: -smp[k])[:top]:
    text = src[ln - 1].strip()[:90] if ln and ln <= len(src) else '?'
    print('%5s  inst %5.1f%%  samples %5.1f%%  wavefronts %9.3g  thr/inst %4.1f | %s' % (ln, 100 * ins[ln] / TI, 100 * smp[ln] / TS, wav[ln], thr[ln] / max(ins[ln], 1), text))
