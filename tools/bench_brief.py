"""Developer tool: the key figures of a bench.py JSON line.  python tools/bench_brief.py FILE"""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("ms/step %.3f  phases %s" % (d["ms_per_step"], {k: round(v, 3) for k, v in d["phases_ms"].items()}))
print("value %.4g edges/s  e2e %.3f ms -> %.4g edges/s  launches %s  clocks %s" % (d["value"], d["e2e"]["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d["clocks"]))
print("roofline: %s frac %.3f | knn frac %.3f executed %.3f | edges frac %.3f" % (d["roofline"]["kernel"], d["roofline"]["frac"], d["roofline_knn"]["frac"], d["roofline_knn"].get("frac_executed", float("nan")), d["roofline_edges"]["frac"]))
if "cpu_baseline" in d:
    print("cpu:", d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"], d["cpu_baseline"].get("python_per_call", {}).get("value"))
for k in d:
    if k.startswith("extra_"):
        print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in d[k].items() if not isinstance(b, (dict, str))})
