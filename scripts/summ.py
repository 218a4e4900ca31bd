import json,sys
d=json.load(open(sys.argv[1]))
bad={k:v for k,v in d.items() if ('mismatch' in k and v) or (k.endswith('equal') and not v) or (k.endswith('bitexact') and not v)}
print(sys.argv[1], 'BAD' if bad else 'ok', bad, {k:(round(v['render_ms'],2), '%.2e'%v['relmse']) for k,v in d.items() if k.startswith('film')})
