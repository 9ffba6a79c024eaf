import json,sys
d=json.load(open(sys.argv[1]))
for k in d:
    v=d[k]
    print(k, round(v.get('ms',0),3), round(v['roofline']['frac'],3) if v.get('roofline') and v['roofline'].get('frac') else None, v.get('kernel_ms') if isinstance(v.get('kernel_ms'),dict) else '')
