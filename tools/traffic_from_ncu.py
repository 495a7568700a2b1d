"""ncu CSV (tools/ncu_tick_traffic.py) + the region's counters -> profiles/r02_traffic.json: measured DRAM bytes per
simulation (search kernel) and per evaluation (network kernel), next to the algorithmic models of DESIGN.md 4."""
import csv
import json
import sys

csv_path, counters_json, out_path = sys.argv[1:4]
rows = list(csv.reader(l for l in open(csv_path) if l.startswith('"')))
hdr = rows[0]
ki, mi, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
idi = hdr.index('ID')
scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1e-9, 'us': 1e-6, 'ms': 1e-3, 'usecond': 1e-6, 'nsecond': 1e-9, 'msecond': 1e-3}
acc = {}
for r in rows[1:]:
    name = r[ki]
    kind = 'search' if 'k_advance' in name else 'net' if 'k_net_forward' in name else None
    if kind is None:
        continue
    a = acc.setdefault(kind, {'ids': set(), 'dram': 0.0, 'time': 0.0})
    a['ids'].add(r[idi])
    v = float(r[vi].replace(',', '')) * scale.get(r[ui], 1.0)
    if r[mi].startswith('dram__bytes'):
        a['dram'] += v
    elif r[mi].startswith('gpu__time_duration'):
        a['time'] += v
c = json.loads(open(counters_json).read().strip().splitlines()[-1])
out = {
    'how': 'ncu dram__bytes_read.sum + dram__bytes_write.sum of every k_advance / k_net_forward_tc launch of one steady-state region of the bench '
           'workload in tick mode (tools/ncu_tick_traffic.py), divided by the simulations / evaluations the engine counted in that region',
    'region': c.get('region'), 'sims': c['sims'], 'evals': c['evals'],
    'search_launches': len(acc['search']['ids']), 'net_launches': len(acc['net']['ids']),
    'search_dram_bytes_per_launch': acc['search']['dram'] / len(acc['search']['ids']),
    'net_dram_bytes_per_launch': acc['net']['dram'] / len(acc['net']['ids']),
    'search_dram_bytes_per_sim': acc['search']['dram'] / c['sims'],
    'net_dram_bytes_per_eval': acc['net']['dram'] / c['evals'],
    'search_us_per_launch_under_ncu': 1e6 * acc['search']['time'] / len(acc['search']['ids']),
    'net_us_per_launch_under_ncu': 1e6 * acc['net']['time'] / len(acc['net']['ids']),
}
json.dump(out, open(out_path, 'w'), indent=1)
print(json.dumps(out))
