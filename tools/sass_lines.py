"""Static SASS instruction count per source line of one kernel (code-size diet).
   python tools/sass_lines.py <object.o> <mangled substring> <source basename> [top]"""
import os, re, subprocess, sys, tempfile
from collections import defaultdict
obj, mangled, srcname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
tmp = tempfile.mkdtemp()
subprocess.check_call(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=tmp, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
dis = subprocess.run(['nvdisasm', '--print-line-info-inline', os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
cnt = defaultdict(int); infn = False; cur = None; pend = []
for l in dis.splitlines():
    if l.startswith('.text.'):
        infn = mangled in l; continue
    if not infn: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m: pend.append((os.path.basename(m.group(1)), int(m.group(2)))); continue
    if re.match(r'\s*/\*[0-9a-f]{4,}\*/\s+\S', l):
        if pend:
            mine = [p for p in pend if p[0] == srcname]
            cur = mine[0][1] if mine else cur
            pend = []
        cnt[cur] += 1
src = open(os.path.join(os.path.dirname(os.path.abspath(obj)), srcname)).read().splitlines()
print('total', sum(cnt.values()))
for ln in sorted(cnt, key=lambda k: -cnt[k])[:top]:
    print('%5s %5d | %s' % (ln, cnt[ln], src[ln - 1].strip()[:100] if ln else '?'))
