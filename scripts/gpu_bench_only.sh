#!/bin/bash
# bench only (no tests), 2 repetitions
for i in 1 2; do timeout 300 python bench.py --no-cpu-baseline --e2e-steps 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['roofline']['frac'],4), d['config']['variant'], d['clocks'])"; done
