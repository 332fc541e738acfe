"""dev helper: summarise an ncu report (raw metrics + per-basic-block instruction counts and stalls)"""
import collections, csv, io, re, subprocess, sys
def opname(s):
    m = re.match(r'(@!?U?P[T\d]+\s+)?([A-Z0-9_.]+)', s.strip())
    return m.group(2).split('.')[0] if m else s.strip()[:12]
rep = sys.argv[1]; W = int(sys.argv[2]) if len(sys.argv) > 2 else 27648
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'lts__t_sector_hit_rate.pct']
for h, u, v in zip(hdr, units, vals):
    if h in want or ('issue_stalled' in h and 'per_issue_active' in h and float(v or 0) > 0.05):
        print(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); hdr = rows[1]
isrc, iex, ith = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Avg. Threads Executed')
st = {k: hdr.index(k) for k in ['stall_long_sb', 'stall_barrier', 'stall_wait', 'stall_short_sb', 'stall_math', 'stall_not_selected', '# Samples']}
data = [r for r in rows[2:] if len(r) > iex]
tot = sum(int(r[iex]) for r in data)
print("total warp inst", tot, "per level-warp", tot / W / 79)
ops = collections.Counter()
for r in data:
    ops[opname(r[isrc])] += int(r[iex])
print("per level-warp by opcode:", {o: round(c / W / 79, 1) for o, c in ops.most_common(22)})
blocks = []; cur = None
for idx, r in enumerate(data):
    e = int(r[iex])
    if cur is None or cur['e'] != e:
        cur = {'e': e, 'start': idx, 'n': 0, 't': float(r[ith]), 'ops': collections.Counter(), 'st': collections.Counter()}; blocks.append(cur)
    cur['n'] += 1
    cur['ops'][opname(r[isrc])] += 1
    for k, i in st.items(): cur['st'][k] += int(r[i])
for b in blocks:
    share = b['e'] * b['n'] / tot
    if share > 0.006:
        s = {k.replace('stall_', ''): v for k, v in b['st'].items() if v > 300}
        print(f"idx {b['start']:5d} n={b['n']:4d} exec/warp={b['e']/W:7.1f} thr={b['t']:4.1f} share={100*share:5.1f}% {dict(b['ops'].most_common(7))} stalls={s}")
