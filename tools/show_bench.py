import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value", round(d["value"]), "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"]), "launches", d["gpu_launches"], "clocks", d["clocks"])
for k,v in d["stages"].items(): print(" ", k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items()})
if d.get("cpu_baseline"): print(" cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
