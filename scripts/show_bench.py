import json,sys
for f in sys.argv[1:]:
    try:
        d=json.load(open(f))
    except Exception as e:
        print(f, "unreadable", e); continue
    r=d["roofline"]
    print(f, "value=%.4g"%d["value"], "e2e=%.4g"%(d["e2e"]["value"] if d.get("e2e") else 0), "ms/step=%.2f"%d["ms_per_step"], "pairs/bead=%.1f"%r["useful_pairs_per_bead"], "frac=%.4f"%r["frac"])
    print("   ", r["step_kernels_ms"])
