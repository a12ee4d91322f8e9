"""One line per captured launch of an ncu report: duration, DRAM bytes, achieved DRAM GB/s, issue slots, active threads per
warp, occupancy, local-memory traffic.  python tools/ncu_table.py <rep.ncu-rep> > profiles/<name>_table.txt"""
import csv, subprocess, sys
txt = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
h = rows[0]
def col(name):
    return h.index(name) if name in h else None
want = [('Kernel Name', 'kernel'), ('gpu__time_duration.sum', 'dur'), ('dram__bytes_read.sum', 'rd'), ('dram__bytes_write.sum', 'wr'),
        ('sm__inst_issued.avg.pct_of_peak_sustained_active', 'issue%'), ('smsp__thread_inst_executed_per_inst_executed.ratio', 'thr/inst'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'), ('smsp__inst_executed_op_local_ld.sum', 'lld'),
        ('smsp__inst_executed_op_local_st.sum', 'lst'), ('launch__registers_per_thread', 'regs')]
idx = [(col(a), b) for a, b in want]
units = rows[1]
def val(r, i):
    if i is None: return float('nan')
    try: return float(r[i].replace(',', ''))
    except Exception: return float('nan')
def scale(i, kind):
    u = units[i] if i is not None else ''
    if kind == 'time': return {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'nsecond': 1e-3, 'usecond': 1.0, 'msecond': 1e3, 'second': 1e6}.get(u, 1.0)
    return {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1.0)
print('%-34s %10s %10s %9s %7s %8s %6s %9s %5s' % ('kernel', 'us', 'dram MB', 'GB/s', 'issue%', 'thr/inst', 'occ%', 'local ld+st', 'regs'))
for r in rows[2:]:
    if len(r) < len(h): continue
    name = r[idx[0][0]].split('(')[0].replace('void ', '')[:34]
    us = val(r, idx[1][0]) * scale(idx[1][0], 'time')
    by = val(r, idx[2][0]) * scale(idx[2][0], 'b') + val(r, idx[3][0]) * scale(idx[3][0], 'b')
    print('%-34s %10.1f %10.1f %9.1f %7.1f %8.1f %6.1f %9.0f %5.0f' % (name, us, by / 1e6, by / 1e3 / us if us > 0 else 0, val(r, idx[4][0]), val(r, idx[5][0]),
                                                                 val(r, idx[6][0]), val(r, idx[7][0]) + val(r, idx[8][0]), val(r, idx[9][0])))
