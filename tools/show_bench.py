import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],2))
print('one_stream',{k:(round(v,3) if isinstance(v,float) else v) for k,v in d['one_stream'].items() if not isinstance(v,(dict,list))})
print(d['one_stream'].get('kernel_ms'))
print(d['roofline'])
