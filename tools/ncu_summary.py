#!/usr/bin/env python
"""Summarise an ncu report (--set full) into profiles/: per-launch key metrics as text + traffic.json.
Usage: python tools/ncu_summary.py gpurun_out/X.ncu-rep profiles/rNN_name.txt "<header comment>" [--traffic]"""
import csv, io, json, os, subprocess, sys

WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_atom.sum', 'sm__cycles_elapsed.avg.per_second']


def main():
    rep, out, header = sys.argv[1], sys.argv[2], sys.argv[3]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines = ['# ' + header, '# source: %s (not committed); values per launch' % rep]
    traffic = {}
    for r in rows[2:]:
        name = r[idx['Kernel Name']]
        lines.append('kernel: ' + name)
        for w in WANT:
            if w in idx and r[idx[w]] != '':
                lines.append('  %-70s %s %s' % (w, r[idx[w]], units[idx[w]]))
        stalls = sorted(((float(r[i] or 0), h) for h, i in idx.items()
                         if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')), reverse=True)
        if stalls:
            lines.append('  top stalls (warps per issue): ' + ', '.join('%s %.2f' % (h.split('stalled_')[1].split('_per_issue')[0], v) for v, h in stalls[:6]))
        def gb(k):
            v, u = float(r[idx[k]]), units[idx[k]]
            return v * {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}.get(u, 1)
        short = name.split('(')[0].replace('void ', '').replace('gsm::', '').split('<')[0].replace('_kernel', '')
        traffic.setdefault(short, []).append(gb('dram__bytes_read.sum') + gb('dram__bytes_write.sum'))
    open(out, 'w').write('\n'.join(lines) + '\n')
    if '--traffic' in sys.argv:
        tj = {k: sum(v) / len(v) for k, v in traffic.items()}
        # stamped with the commit the capture was taken from (commit BEFORE the GPU visit): bench.py reports it next to
        # roofline.traffic, so a figure from an older build is visible as such
        head = subprocess.run(['git', 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True).stdout.strip()
        tj['_commit'] = head or os.environ.get('GSM_COMMIT', '')   # (no .git on the GPU box: the visit script passes it)
        tj['_source'] = os.path.basename(out)
        json.dump(tj, open(os.path.join(os.path.dirname(out), 'traffic.json'), 'w'), indent=1)
    print('wrote', out)


if __name__ == '__main__':
    main()
