"""Summarise an RFK_PAIR_TRACE file (development): per-role intervals of pair 0, both CTAs."""
import collections
import statistics
import sys


def med(v):
    return 'n=%d med %.0f mean %.0f' % (len(v), statistics.median(v), statistics.mean(v)) if v else 'n=0'


def main(path):
    ev = collections.defaultdict(list)
    for line in open(path):
        c, role, _, t, tag = line.split()
        ev[(int(c), role)].append((int(t), int(tag)))
    for c in (0, 1):
        print('--- CTA', c)
        items, cur = [], {}
        for t, tag in ev[(c, 'epi')]:
            if tag == 1 and cur:
                items.append(cur)
                cur = {}
            cur[tag] = t
        if cur:
            items.append(cur)
        I = items[2:-1]
        print(' items', len(items), ' item period', med([items[k + 1][1] - items[k][1] for k in range(2, len(items) - 1)]))
        for a, b, nm in ((1, 5, 'acc1 seen -> acc3 seen'), (5, 2, 'acc3 seen -> epi1 done'), (2, 6, 'epi1 done -> softmax done'),
                         (6, 3, 'softmax done -> acc2 seen'), (3, 4, 'acc2 seen -> epi2 done')):
            print('  %-28s' % nm, med([i[b] - i[a] for i in I if a in i and b in i]))
        print('  %-28s' % 'epi2 done -> next acc1', med([items[k + 1][1] - items[k][4] for k in range(2, len(items) - 1) if 4 in items[k]]))
        for role, tags in (('conv', {(1, 2): 'raw seen -> a_free seen', (2, 3): 'a_free seen -> arrived', (1, 1): 'raw-stage period'}),
                           ('mma', {(1, 2): 'a_full seen -> issued', (2, 1): 'issued -> next a_full seen', (1, 1): 'atom period'}),
                           ('tail', {(1, 2): 'hid(L2) seen -> L2 committed', (2, 3): 'L2 committed -> hid(L3) seen',
                                     (3, 4): 'hid(L3) seen -> L3 committed', (4, 1): 'L3 committed -> next hid(L2)'})):
            last, per = {}, collections.defaultdict(list)
            for t, tag in ev[(c, role)]:
                for (a, b), nm in tags.items():
                    if tag == b and a in last:
                        per[nm].append(t - last[a])
                last[tag] = t
            for nm, v in per.items():
                print('  %s: %-30s' % (role, nm), med(v[10:]))


if __name__ == '__main__':
    main(sys.argv[1])
