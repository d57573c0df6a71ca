import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    if "H" in d: print(d["H"],d["C"],{k:v for k,v in d.items() if k.endswith("GBs")})
    else: print(d)
