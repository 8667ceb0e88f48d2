#!/usr/bin/env python
"""Digest an `ncu --page source --csv` export: opcode mix by executed warp-instructions and the
execution-count bands of the SASS listing (which loop bodies cost what).
Usage: python tools/sass_profile.py source.csv [--bands]"""
import collections, csv, sys

def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = next(r for r in rows if 'Instructions Executed' in r)
    ix = {h: i for i, h in enumerate(hdr)}
    tot = 0; ops = collections.Counter(); data = []; seen_hdr = 0
    for r in rows:
        if 'Instructions Executed' in r:
            seen_hdr += 1
            if seen_hdr > 1: break
            continue
        if len(r) <= ix['Instructions Executed'] or not r[ix['Instructions Executed']].isdigit(): continue
        n = int(r[ix['Instructions Executed']]); tot += n
        toks = r[ix['Source']].split()
        op = toks[1] if toks[0].startswith('@') else toks[0]
        ops[op.split('.')[0]] += n
        data.append((n, float(r[ix['Avg. Threads Executed']] or 0), r[ix['Source']].strip(), int(r[ix['# Samples']] or 0)))
    print('SASS instructions', len(data), 'executed warp-instructions', tot, 'samples', sum(d[3] for d in data))
    print(' '.join(f'{k} {v / tot * 100:.1f}%' for k, v in ops.most_common(22)))
    if '--bands' in sys.argv:
        prev = None; acc_n = 0; acc_s = 0; start = 0
        def flush(i):
            if prev is not None:
                print(f'  [{start:4d}-{i - 1:4d}] {i - start:4d} instr x {prev:8.2f}M  = {acc_n / 1e6:8.1f}M ({acc_n / tot * 100:4.1f}%) samples {acc_s / max(1, sum(d[3] for d in data)) * 100:4.1f}%   {data[start][2][:50]}')
        for i, (n, t, src, smp) in enumerate(data):
            b = n / 1e6
            if prev is None or abs(b - prev) > 0.2 * max(b, prev, 0.3):
                flush(i); prev = b; acc_n = 0; acc_s = 0; start = i
            acc_n += n; acc_s += smp
        flush(len(data))

if __name__ == '__main__':
    main()
