import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f)); r=d["roofline"]
        print(f, "ms", round(d["ms_per_step"],4), "kernel_ms", round(r["kernel_ms"],4), "frac", round(r["frac"],4), "frac_ref", round(r.get("frac_of_reference_work",0),4), "e2e_ms", round(d["e2e"]["ms_per_step"],3), "val %.4g"%d["value"], "e2e_free_ms", (d.get("e2e_free_parameters") or {}).get("ms_per_step"))
    except Exception as e: print(f, "ERR", e)
