import json,sys
line=[l for l in open(sys.argv[1]) if l.startswith('{')][-1]
d=json.loads(line)
print('value',d['value'],'ms/step',d['ms_per_step'],'frac',d['roofline']['frac'],'clk',d.get('clocks'))
for r in d.get('per_shape',[]): print(' ',r['shape'],'M',r['M'],r['us'],'us',r['GBps'],'GB/s',r['frac_measured_hbm'])
if 'prefill' in d:
    print('prefill M4096', d['prefill']['M4096_TFLOPS'])
    for r in d['prefill']['points']: print(' ',r['shape'],'M',r['M'],r['ms'],'ms',r['TFLOPS'],'TF',r['kernel'])
if 'e2e' in d: print('e2e',d['e2e'])
