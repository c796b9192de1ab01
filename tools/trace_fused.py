"""Print a merged timeline of the in-kernel event trace written with RFK_FUSED_PROF=2 (development tool)."""
import collections
import sys

NAMES = {('conv', 1): 'conv: waits done', ('conv', 2): 'conv: stored', ('conv', 3): 'conv: arrived',
         ('load', 1): 'load: raw TMA issued', ('load', 3): 'load: W1 bulk issued',
         ('mma', 1): 'MMA: operands ready', ('mma', 2): 'MMA: issued', ('mma', 3): 'MMA: committed',
         ('epi', 1): 'EPI: acc1_full seen', ('epi', 2): 'EPI: epi1 done', ('epi', 3): 'EPI: acc2_full seen',
         ('epi', 4): 'EPI: epi2 done', ('epi', 5): 'EPI: acc3 seen', ('epi', 6): 'EPI: softmax done',
         ('tail', 1): 'TAIL: hid(L2) seen', ('tail', 2): 'TAIL: L2 issued', ('tail', 3): 'TAIL: hid(L3) seen',
         ('tail', 4): 'TAIL: L3 issued'}
INDENT = {'conv': '', 'load': '    ', 'mma': '        ', 'epi': '                ', 'tail': '                    '}


def main(path, first_mma_event=600, n_events=70):
    ev = collections.defaultdict(list)
    for line in open(path):
        role, _, t, tag = line.split()
        ev[role].append((int(t), int(tag)))
    t0 = min(e[0][0] for e in ev.values())
    allev = sorted((t - t0, r, tag) for r, l in ev.items() for t, tag in l)
    m = ev['mma']
    start = m[first_mma_event][0] - t0
    end = m[min(first_mma_event + n_events, len(m) - 1)][0] - t0
    prev = None
    for t, r, tag in allev:
        if start <= t <= end:
            print('%8d %+6d  %s%s' % (t, 0 if prev is None else t - prev, INDENT[r], NAMES[(r, tag)]))
            prev = t
    # per-step statistics of the layer-1 issuer
    import statistics
    per = collections.defaultdict(list)
    last = {}
    for t, tag in m:
        if tag == 1 and 3 in last: per['commit->ready'].append(t - last[3])
        if tag == 2 and 1 in last: per['ready->issued'].append(t - last[1])
        if tag == 3 and 2 in last: per['issued->committed'].append(t - last[2])
        if tag == 1 and 1 in last: per['step period'].append(t - last[1])
        last[tag] = t
    for k, v in per.items():
        v = v[50:]
        print('%-20s median %6.0f mean %6.0f' % (k, statistics.median(v), statistics.mean(v)))
    c = ev['conv']
    per = collections.defaultdict(list)
    last = {}
    for t, tag in c:
        if tag == 2 and 1 in last: per['conv waits->stored'].append(t - last[1])
        if tag == 3 and 2 in last: per['conv stored->arrived'].append(t - last[2])
        if tag == 1 and 3 in last: per['conv arrived->next waits done'].append(t - last[3])
        last[tag] = t
    for k, v in per.items():
        v = v[50:]
        print('%-30s median %6.0f mean %6.0f' % (k, statistics.median(v), statistics.mean(v)))


if __name__ == '__main__':
    main(sys.argv[1], *[int(a) for a in sys.argv[2:]])
