import json, sys
txt = open(sys.argv[1]).read().strip().splitlines()
d = json.loads([l for l in txt if l.startswith("{")][-1])
print(sys.argv[2], "N", d["n_gpus"], "img/s", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 2),
      {k: d["optimizer_step"][k] for k in ("us", "fused")}, "dot us", round(d["roofline"]["us_per_launch"], 1),
      "GB/s", round(d["roofline"]["achieved"]), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
