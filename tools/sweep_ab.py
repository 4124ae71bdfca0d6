import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    w=d["config"]["workload"]
    if w.endswith(("_C1","_C4","_C16")): print(w, round(d["ms_apply"],4), round(d["roofline"]["frac"],3))
