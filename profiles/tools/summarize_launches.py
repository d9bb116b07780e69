"""Per-kernel totals of an ncu launch list (profiles/r1_launches.csv):
    python profiles/tools/summarize_launches.py profiles/r1_launches.csv > profiles/r1_launches_summary.txt"""
import collections
import csv
import sys

HEAD = """ncu launch list of `python bench.py --steps 20 --warmup 5 --no-cpu --no-extras` (B200; `ncu --metrics
gpu__time_duration.sum --clock-control none`, first 600 launches; cold-cache serialised times: compare shares, not
absolutes).  The timed region of the bench is the 20 match_count launches of the step loop (one launch per step);
the other launches are warm-up, the oracle slice check, the end-to-end legs (12 chunk launches per step; the raw
form adds one pack_heads launch per chunk), and the replay leg (default policy, sweep, int32 and uint16 traces).
"""


def main(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ik, iv, iu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.OrderedDict()
    match = []
    for r in rows[1:]:
        if len(r) <= iv or not r[iv]:
            continue
        ms = float(r[iv].replace(',', '')) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0}[r[iu]]
        a = agg.setdefault(r[ik], [0, 0.0])
        a[0] += 1
        a[1] += ms
        if 'match_count' in r[ik]:
            match.append(ms)
    tot = sum(a[1] for a in agg.values())
    print(HEAD)
    for k, (n, t) in agg.items():
        print('%-92s n=%3d total=%9.3f ms  share=%5.1f%%  avg=%.4f ms' % (k[:92], n, t, 100 * t / tot, t / n))
    full = [m for m in match if m > 0.4]
    print()
    print('match_count launches over the full 1e8-key snapshot: n=%d avg=%.4f ms' % (len(full), sum(full) / len(full)))


if __name__ == '__main__':
    main(sys.argv[1])
