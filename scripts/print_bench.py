import json,sys
for l in sys.stdin.read().strip().splitlines()[-1:]:
    d=json.loads(l); print(sys.argv[1], round(d["value"],1), {k:round(v/9,3) for k,v in d["phase_ms_per_step"].items()}, (d.get("parity") or {}).get("rel_diff"), "e2e", d.get("e2e") and round(d["e2e"]["value"],1), "ptr", d.get("e2e_ptr") and round(d["e2e_ptr"]["value"],1))
