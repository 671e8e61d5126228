#!/usr/bin/env python
"""Print the e2e block and the e2e_sweep table of a bench.py JSON line."""
import json, sys
d = json.load(open(sys.argv[1]))
e = d['e2e']
print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in e.items() if k != 'best_case'})
print(e.get('best_case'))
sw = d.get('e2e_sweep')
if sw:
    print('probe', sw['pcie_probe_gbs_each_way'])
    for blk, rows in sw['blocks'].items():
        print(blk)
        for k, row in rows.items():
            print('  %-20s' % k, '  '.join('%s:%s' % (a, (b or {}).get('msps', (b or {}).get('error', '-'))) for a, b in row.items()))
