"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.  usage: python profiles/launch_summary.py launches.csv"""
import collections
import csv
import sys


def main(path):
    lines = [l for l in open(path) if not l.startswith('==')]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        k = row['Kernel Name'].split('(')[0].replace('void ', '').replace('gsr::<unnamed>::', '')
        v = float(row['Metric Value'].replace(',', ''))
        u = row['Metric Unit']
        v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
        agg.setdefault(k, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print(f"launches {sum(len(v) for v in agg.values())}, total {tot:.1f} us (cold-cache, serialised: compare shares)")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k[:64]:64s} n={len(v):4d} sum={sum(v):10.1f} us avg={sum(v) / len(v):9.1f} us share={sum(v) / tot:.3f}")


if __name__ == '__main__':
    main(sys.argv[1])
