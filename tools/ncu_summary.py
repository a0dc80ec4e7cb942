"""Summarise an ncu report: per-launch headline metrics (raw page) and, for one kernel,
the instruction/stall distribution by SASS block (source page)."""
import csv, subprocess, sys

KEYS = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'lts__t_sector_hit_rate.pct', 'sm__cycles_elapsed.avg',
        'smsp__warps_eligible.avg.per_cycle_active']


def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    keys = KEYS + [h for h in hdr if 'issue_stalled' in h and 'per_issue_active' in h]
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            vals = [r[i][:24] for r in rows[2:]]
            if any(v not in ('0', '0.000000', '-nan') for v in vals):
                print(f"{k[:78]:78s} | " + " | ".join(vals))


def source(rep, kernel, launch=0, lo=None, hi=None):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kernel}',
                          '--launch-skip', str(launch), '--launch-count', '1'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ia, isrc, ie, iss = (hdr.index(x) for x in ('Address', 'Source', 'Instructions Executed', '# Samples'))
    seen, data = set(), []
    for r in rows[2:]:
        if len(r) <= ie or r[ia] in seen:
            continue
        seen.add(r[ia])
        try:
            data.append((int(r[ia], 16), r[isrc], int(r[ie]), int(r[iss])))
        except ValueError:
            pass
    base = data[0][0]
    tot = sum(d[2] for d in data)
    ts = sum(d[3] for d in data)
    print(rows[0][1], "instructions", tot, "stall samples", ts)
    if lo is not None:
        for d in data:
            off = d[0] - base
            if lo <= off <= hi:
                print(f"{off:6x} {d[2]:12d} {d[3]:7d}  {d[1][:100]}")
        return
    blk, cur = [], None
    for d in data:
        if cur and abs(d[2] - cur[2]) <= 0.02*max(cur[2], 1):
            cur[1] += 1; cur[3] += d[2]; cur[4] += d[3]
        else:
            if cur:
                blk.append(cur)
            cur = [d[0] - base, 1, d[2], d[2], d[3]]
    blk.append(cur)
    for b in blk:
        if b[3] > 0.003*tot or b[4] > 0.003*ts:
            print(f"{b[0]:6x} n={b[1]:4d} exec_each={b[2]:12d} inst_share={b[3]/tot:.3f} stall_share={b[4]/ts:.3f}")


if __name__ == '__main__':
    rep = sys.argv[1]
    if len(sys.argv) == 2:
        raw(rep)
    else:
        a = sys.argv[2:]
        source(rep, a[0], int(a[1]) if len(a) > 1 else 0, int(a[2], 16) if len(a) > 2 else None,
               int(a[3], 16) if len(a) > 3 else None)


def ranges(rep, kernel, launch, cuts):
    """Instruction / stall shares between address cut points (hex offsets)."""
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', f'regex:{kernel}',
                          '--launch-skip', str(launch), '--launch-count', '1'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[1]
    ia, ie, iss = (hdr.index(x) for x in ('Address', 'Instructions Executed', '# Samples'))
    seen, data = set(), []
    for r in rows[2:]:
        if len(r) <= ie or r[ia] in seen:
            continue
        seen.add(r[ia])
        try:
            data.append((int(r[ia], 16), int(r[ie]), int(r[iss])))
        except ValueError:
            pass
    base = data[0][0]
    tot = sum(d[1] for d in data); ts = sum(d[2] for d in data)
    cuts = [0] + cuts + [1 << 30]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        si = sum(d[1] for d in data if lo <= d[0] - base < hi)
        ss = sum(d[2] for d in data if lo <= d[0] - base < hi)
        mx = max([d[1] for d in data if lo <= d[0] - base < hi] or [0])
        print(f"[{lo:6x},{hi:9x}) inst={si/tot:.3f} stall={ss/ts:.3f} max_exec={mx:.3e}")
