#!/usr/bin/env python
"""pretty-print a bench.py JSON line"""
import json
import sys

d = json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith('{')][-1])
print(f"value {d['value']:.0f} {d['unit']}  ms/step {d['ms_per_step']:.3f}  e2e {d['e2e']['value']:.0f}  launches {d.get('gpu_launches')}"
      f"  clocks {d.get('clocks')}")
print('roofline', d.get('roofline'))
print('mlp', d.get('mlp'))
for k, v in d.get('stages', {}).items():
    print(f'  {k:26s}', v)
if 'cpu_baseline' in d:
    print('cpu', d['cpu_baseline'])
