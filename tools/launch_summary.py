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


def level_table(seq):
    """One boosting round (from k_grad to k_round_end) as a kernel x tree-level table of durations in us."""
    g = [i for i, s in enumerate(seq) if s[0] == 'k_grad'][0]
    rnd = seq[g:]
    rnd = rnd[:[i for i, s in enumerate(rnd) if s[0] == 'k_round_end'][0] + 1]
    lvl, table = -1, collections.defaultdict(dict)
    for k, v, _, _ in rnd:
        if k == 'k_plan_hist':
            lvl += 1
        if k in ('k_grad', 'k_predict', 'k_round_end'):
            table[k]['all'] = v
        else:
            table[k][lvl] = table[k].get(lvl, 0) + v
    lines = ['| kernel | ' + ' | '.join('L%d' % l for l in range(lvl + 1)) + ' | sum |', '|---|' + '---|' * (lvl + 2)]
    for k in ('k_hist', 'k_scan', 'k_scan_deep', 'k_part_count', 'k_part_scatter', 'k_hist_zero', 'k_plan_hist', 'k_plan_part'):
        row = [table[k].get(l) for l in range(lvl + 1)]
        lines.append('| %s | ' % k + ' | '.join('%.0f' % v if v is not None else '' for v in row) + ' | %.0f |' % sum(v for v in row if v))
    lines.append('')
    lines.append('Per round: k_grad %.0f, k_predict %.0f, k_round_end %.0f us; whole round %.1f ms.'
                 % (table['k_grad']['all'], table['k_predict']['all'], table['k_round_end']['all'], sum(v for _, v, _, _ in rnd) / 1000))
    return '\n'.join(lines)
