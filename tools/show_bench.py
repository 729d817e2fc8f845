"""python tools/show_bench.py FILE... : one-screen summary of bench.py JSON lines."""
import json, sys
for f in sys.argv[1:]:
    try:
        l = [x for x in open(f) if x.startswith("{")]
        d = json.loads(l[-1])
    except Exception as ex:
        print(f, "no line:", ex); continue
    print(f, "value", round(d["value"]), "ms", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["value"]),
          "launches/step", d["gpu_launches"] / d["steps"], "roofline", round(d["roofline"]["frac"], 4))
    print("  variants:", d["config"].get("kernel_variants"))
    for k in ("greedy", "greedy_large", "c4", "device_dataset", "greedy_sharded"):
        v = d.get(k)
        if v:
            print("  ", k, {a: (round(b, 1) if isinstance(b, float) else b) for a, b in v.items() if a in ("value", "ms_per_step", "ms_per_batch", "error")},
                  (v.get("b16384") or {}).get("value"))
    for k in d["kernels"][:30]:
        print(f"   {k['ms']*1e3:8.1f} us  {k['name'][:100]}")
