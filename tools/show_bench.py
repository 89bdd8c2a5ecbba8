"""Print the interesting fields of a bench.py JSON line."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for k in ('value', 'ms_per_step', 'e2e', 'gpu_launches', 'clocks', 'cpu_baseline', 'proves_per_sec_n64_m16', 'verifies_per_sec_per_proof_challenges',
          'verifies_per_sec_rlc_batch', 'proves_per_sec_interactive', 'msm_point_range_split'):
    print(k, d.get(k))
r = d['roofline']
print({k: r.get(k) for k in ('achieved', 'peak', 'frac', 'avg_launch_ms', 'kernel_share_of_step', 'whole_step')})
for s in d.get('msm_points_per_sec', []):
    print(s)
