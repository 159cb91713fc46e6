import json, sys
d = json.load(open(sys.argv[1]))
print('value %.1f GB/s  ms/step %.3f  frac %.3f  launches/step %s clocks %s' % (d['value'], d['ms_per_step'], d['frac_hbm_measured'], d['launches_per_step'], d['clocks']))
print('e2e %.1f GB/s ms/step %.2f' % (d['e2e']['value'], d['e2e']['ms_per_step']))
for c in d['cases']:
    ch = c.get('chain') or {}
    print('  K=%5d N=%5d M=%2d  %7.2f us/layer  %7.1f GB/s  frac %.3f  tok/s %.0f | chain %.2f us/layer %.0f GB/s' % (c['K'], c['N'], c['M'], c['us_per_layer'], c['gbs'], c['frac_hbm_measured'], c['tokens_per_s'], ch.get('us_per_layer', 0), ch.get('gbs', 0)))
if d.get('gemm'):
    g = d['gemm']; print('gemm %.1f TOPS (%.3f of 4.5P) ms %.4f  incl quant %.1f TOPS' % (g['tops'], g['roofline']['frac'], g['ms_gemm'], g['tops_incl_act_quant']))
if d.get('cpu_baseline'): print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'])
if d.get('vit_l14'):
    for v in d['vit_l14']: print('  vit K=%d N=%d M=%d %.1f us %.0f TOPS' % (v['K'], v['N'], v['M'], v['us'], v['tops_incl_act_quant']))
