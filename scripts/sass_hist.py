#!/usr/bin/env python
"""static opcode histogram of a SASS dump (cuobjdump -sass -fun ...): python scripts/sass_hist.py file.sass [samples_per_iteration]"""
import collections
import re
import sys

ops = collections.Counter(); n = 0
per = float(sys.argv[2]) if len(sys.argv) > 2 else 16.0
for l in open(sys.argv[1]):
    m = re.match(r'\s+/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)', l)
    if m:
        ops['.'.join(m.group(2).split('.')[:2])] += 1; n += 1
print(n, n / per)
for k, v in ops.most_common(50):
    print(f'{k:18s}{v:5d} {v / per:6.2f}')
