import json, sys
d = json.load(open(sys.argv[1]))
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches", "n_gpus")}, "e2e", round(d["e2e"]["value"], 1),
      "roofline", round(d["roofline"]["achieved"]), round(d["roofline"]["frac"], 3), d["clocks"])
for k, v in d.get("kernels", {}).items():
    print(f"{k:32s} {v['us']:8.1f} us {v['GB/s']:8.0f} GB/s  of_meas={v['frac_of_measured_peak']:.3f} of8={v.get('frac_of_8TBs', 0):.3f}")
print(d.get("cpu_baseline"))
