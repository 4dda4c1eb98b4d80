import json,sys
for l in sys.stdin.read().strip().splitlines()[-1:]:
    d=json.loads(l); print(sys.argv[1], round(d["value"],1), {k:round(v/9,3) for k,v in d["phase_ms_per_step"].items()}, d["parity"]["rel_diff"])
