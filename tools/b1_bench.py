import json, sys, torch
sys.path.insert(0, '.')
import bench
for r in bench.run_binary_1bit(torch.device('cuda:0'), bench.peaks()):
    print(r['M'], {k: (round(v['us_per_layer'], 2), round(v['gbs'])) for k, v in r.items() if isinstance(v, dict)}, round(r['speedup_one_bit'], 2))
