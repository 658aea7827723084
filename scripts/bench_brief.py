import json, sys
d = json.load(open(sys.argv[1]))
print("value %.1f G evals/s  ms/step %.2f  phases %s" % (d["value"]/1e9, d["ms_per_step"], {k: round(v,3) for k,v in d["phases_ms_per_step"].items()}))
print("rechecked %.4f  roofline frac %.3f  kernel_ms %.2f  learn_wall %.3f s  e2e %.1f G evals/s (%.1f ms)" % (
    d["rechecked_frac"], d["roofline"]["frac"], d["roofline"]["kernel_ms"], d["learn_wall_s"], d["e2e"]["value"]/1e9, d["e2e"]["ms_per_step"]))
print("stages ms:", {k: round(v,1) for k,v in d["learn_stage_ms"].items()})
print("cpu:", d.get("cpu_baseline"), "clocks:", d.get("clocks"))
