import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1] if len(sys.argv) > 1 else "", "cold %.1f warm %.1f e2e %.1f launches/step %d" % (d["value"], d["value_warm_l2"], d["e2e"]["value"], d["gpu_launches"] // d["steps"]),
      " ".join(f"{k}={v*1e3:.1f}" for k, v in d["stage_ms"].items()))
