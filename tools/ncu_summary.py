"""Summarise gpurun_out/*.ncu-rep (ncu --set full captures) and the launch list into profiles/ (tracked).

  python tools/ncu_summary.py r1b        # writes profiles/<tag>_ncu_<kernel>.csv, <tag>_ncu_traffic.json,
                                         #        profiles/<tag>_launches.csv and <tag>_launches_summary.csv
bench.py reads the newest *_ncu_traffic.json for the `roofline.traffic` field (DRAM bytes per launch)."""
import csv
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_tensor', 'sm__pipe_tensor', 'sm__pipe_alu_cycles_active.avg',
        'sm__pipe_fma_cycles_active.avg', 'sm__pipe_fmaheavy_cycles_active.avg', 'sm__inst_executed_pipe_uniform',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block',
        'launch__occupancy_limit', 'sm__cycles_active.avg', 'sm__cycles_elapsed.max', 'smsp__cycles_active.avg',
        'smsp__average_warp_latency_issue_stalled', 'smsp__average_warps_issue_stalled', 'l1tex__data_bank_conflicts',
        'smsp__inst_executed_op_shared', 'smsp__sass_inst_executed_op_shared', 'sm__sass_inst_executed_op_shared',
        'smsp__pcsamp_warps_issue_stalled')


def rows_of(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else 'r1'
    src = os.path.join(ROOT, 'gpurun_out')
    dst = os.path.join(ROOT, 'profiles')
    traffic = {}
    for rep in sorted(glob.glob(os.path.join(src, 'full_*.ncu-rep'))):
        r = rows_of(rep)
        if len(r) < 3:
            continue
        h, u, v = r[0], r[1], r[2]
        name = os.path.basename(rep)[5:-8]
        d = dict(zip(h, v))
        with open(os.path.join(dst, f'{tag}_ncu_{name}.csv'), 'w', newline='') as f:
            w = csv.writer(f)
            w.writerow(['metric', 'unit', 'value'])
            w.writerow(['Kernel Name', '', d.get('Kernel Name', name)])
            for i, n in enumerate(h):
                if any(k in n for k in KEEP) and v[i] not in ('', 'n/a'):
                    w.writerow([n, u[i], v[i]])
        ub = dict(zip(h, u))
        scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}

        def b(k):
            return float(d[k].replace(',', '')) * scale.get(ub[k], 1)
        dur_unit = {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0}[ub['gpu__time_duration.sum']]
        traffic[name] = {'dram_read_bytes': b('dram__bytes_read.sum'), 'dram_write_bytes': b('dram__bytes_write.sum'),
                         'duration_ms_under_ncu': float(d['gpu__time_duration.sum'].replace(',', '')) * dur_unit}
    if traffic:
        json.dump(traffic, open(os.path.join(dst, f'{tag}_ncu_traffic.json'), 'w'), indent=1)
    # source-level view of the sampling kernel: address, samples, executions, dominant stall, SASS
    rep = os.path.join(src, 'full_fps_tmem_kernel.ncu-rep')
    if os.path.exists(rep):
        out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
        rows = list(csv.reader(out.splitlines()))
        if len(rows) > 2:
            hdr = rows[1]
            ci = {h: i for i, h in enumerate(hdr)}
            stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
            with open(os.path.join(dst, f'{tag}_ncu_fps_tmem_kernel_source.csv'), 'w', newline='') as f:
                w = csv.writer(f)
                w.writerow(['address', 'samples', 'instructions_executed', 'top_stall', 'sass'])
                for r in rows[2:]:
                    n = int(r[ci['# Samples']] or 0)
                    top = max(stalls, key=lambda h: int(r[ci[h]] or 0)) if n else ''
                    w.writerow([r[ci['Address']][-5:], n, r[ci['Instructions Executed']], top[6:], r[ci['Source']]])
    lp = os.path.join(src, 'launches.csv')
    if os.path.exists(lp):
        lines = [l for l in open(lp) if l.startswith('"')]
        open(os.path.join(dst, f'{tag}_launches.csv'), 'w').writelines(lines)
        agg = {}
        for row in csv.DictReader(lines):
            if row.get('Metric Name') != 'gpu__time_duration.sum':
                continue
            t = float(row['Metric Value'].replace(',', ''))
            t *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3}.get(row['Metric Unit'], 1.0)
            k = row['Kernel Name'].split('(')[0]
            a = agg.setdefault(k, [0, 0.0])
            a[0] += 1
            a[1] += t
        tot = sum(a[1] for a in agg.values())
        with open(os.path.join(dst, f'{tag}_launches_summary.csv'), 'w', newline='') as f:
            w = csv.writer(f)
            w.writerow(['kernel', 'launches', 'total_us', 'avg_us', 'share'])
            for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                w.writerow([k, a[0], f'{a[1]:.1f}', f'{a[1] / a[0]:.2f}', f'{a[1] / tot:.4f}'])
    for n in ('bench.json', 'bench_ref.json'):
        p = os.path.join(src, n)
        if os.path.exists(p) and os.path.getsize(p):
            open(os.path.join(dst, f'{tag}_{n}'), 'w').write(open(p).read())


if __name__ == '__main__':
    main()
