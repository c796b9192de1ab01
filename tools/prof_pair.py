"""Summarise RFK_PAIR_PROF files (development): cycles per loop segment of each role, pair 0."""
import sys

NAMES = {'epi': ['ovh', 'acc1 wait', 'acc3 fetch', 'epi1', 'fin:rest', 'acc2 wait', 'epi2', 'store-side', 'fin:math|wait full', 'fin:free wait|store', 'fin:sts', 'x'],
         'conv': ['ovh', 'raw wait', 'lds+cvt', 'a_free wait', 'sttm+wait', 'arrive', '-', 'end'],
         'load': ['issue', 'raw_empty wait', '', '', '', '', '', 'end'],
         'mma': ['ovh', 'acc1_empty wait', 'a_full wait', 'issue', 'commit', '', '', 'end'],
         'tail': ['ovh', 'hid wait L2', 'L2 issue', 'hid wait L3', 'L3 issue', '', '', 'end']}
for f in sys.argv[1:]:
    print(f.split('/')[-1])
    for line in open(f):
        v = line.split()
        c, w = int(v[0]), int(v[1])
        a = [int(x) for x in v[2:]]
        role = 'epi' if w < 8 else 'conv' if w < 16 else {16: 'load', 17: 'mma', 18: 'fwd', 19: 'tail'}[w]
        if role == 'fwd' or (c == 1 and role in ('mma', 'tail')) or sum(a) == 0:
            continue
        if (role == 'epi' and w not in (0, 4)) or (role == 'conv' and w not in (8, 12)):
            continue
        print('  cta%d w%2d %-4s tot %7d: ' % (c, w, role, sum(a)) +
              ', '.join('%s %d' % (n, x) for n, x in zip(NAMES[role], a) if n and n != '-'))
