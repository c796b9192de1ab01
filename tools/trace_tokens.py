"""Token timeline of the pair kernel's A ring from an RFK_PAIR_TRACE file (development): for every K atom,
when the issuer saw it full, when its MMAs were issued, when the stage came back (a_free observed by an idle
warp), and what the converter warp 8 of CTA 0 was doing."""
import collections
import statistics
import sys

ev = collections.defaultdict(list)
for line in open(sys.argv[1]):
    c, role, _, t, tag = line.split()
    ev[(int(c), role)].append((int(t), int(tag)))
t0 = min(t for v in ev.values() for t, _ in v)
seen = [t - t0 for t, tag in ev[(0, 'mma')] if tag == 1]
issued = [t - t0 for t, tag in ev[(0, 'mma')] if tag == 2]
freed = [t - t0 for t, tag in ev[(0, 'load')] if tag == 2]
n = min(len(seen), len(issued), len(freed))
print('atoms', n)
lat = [freed[i] - issued[i] for i in range(20, n)]
print('issued -> a_free visible: median %d  p10 %d  p90 %d' % (statistics.median(lat), sorted(lat)[len(lat) // 10],
                                                              sorted(lat)[9 * len(lat) // 10]))
per = [seen[i + 1] - seen[i] for i in range(20, n - 1)]
print('a_full seen period: median %d mean %d' % (statistics.median(per), statistics.mean(per)))
conv = collections.defaultdict(list)
for c in (0, 1):
    cur = {}
    for t, tag in ev[(c, 'conv')]:
        cur[tag] = t - t0
        if tag == 3:
            conv[c].append(cur)
            cur = {}
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 100
for i in range(lo, lo + 30):
    print('atom %4d: seen %7d issued %7d (+%4d) freed %7d (+%4d after issue)' % (i, seen[i], issued[i], issued[i] - seen[i],
                                                                                 freed[i], freed[i] - issued[i]))
print('converter warp 8 of cta0 (raw seen, a_free seen, arrived):')
for it in conv[0][lo // 4: lo // 4 + 10]:
    print('   ', it.get(1), it.get(2), it.get(3), ' cvt+sttm+arrive %d' % (it.get(3, 0) - it.get(2, 0)))
