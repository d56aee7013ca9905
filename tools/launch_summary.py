"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list: per-kernel totals and the launch sequence."""
import collections, csv, re, sys

def load(path):
    lines = [l for l in open(path) if not l.startswith('==')]
    seq = []
    for row in csv.DictReader(lines):
        if row.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        k = re.sub(r'\(.*', '', row['Kernel Name']).split('::')[-1].split('<')[0]
        v = float(row['Metric Value'].replace(',', ''))
        u = row['Metric Unit']
        v = v / 1000 if u == 'ns' else v * 1000 if u == 'ms' else v
        seq.append((k, v, row['Grid Size'], row['Block Size']))
    return seq

if __name__ == '__main__':
    seq = load(sys.argv[1])
    tot, n = collections.Counter(), collections.Counter()
    for k, v, g, b in seq:
        tot[k] += v; n[k] += 1
    T = sum(tot.values())
    print('| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|')
    for k, v in tot.most_common():
        print('| %s | %d | %.1f | %.1f | %.1f%% |' % (k, n[k], v, v / n[k], 100 * v / T))
    if len(sys.argv) > 2:
        for s in seq:
            print('%-16s %9.1f  %s %s' % s)
