import json,sys
d=json.load(open(sys.argv[1]))
def show(x):
    print(x['metric'], '%.3fM'%(x['value']/1e6), 'ms %.1f'%x['ms_per_step'], 'e2e %.3fM'%(x['e2e']['value']/1e6), 'roof %.3f share %.3f kms %.1f'%(x['roofline']['frac'], x['roofline']['kernel_share_of_step'], x['roofline']['kernel_ms_per_step']), x['clocks'], x['gpu_launches'])
show(d)
if 'also' in d: show(d['also'])
print(d.get('cpu_baseline'))
