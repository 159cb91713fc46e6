import json, sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '.')
import bench
dev = torch.device('cuda:0')
for r in bench.run_prep(dev, bench.peaks()):
    print(r['kernel'][:64].ljust(64), round(r['us'], 1), 'us', round(r['gbs']), 'GB/s', round(r['frac_hbm_measured'], 3))
