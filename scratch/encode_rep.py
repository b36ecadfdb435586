# repeat the c3x whole-image encode leg of bench.py a few times to see the run-to-run spread of each phase, with the fixed
# costs of train() (initial test, trainer construction, validations, final pull) timed separately
import sys, time
sys.path.insert(0, ".")
import torch
import bench
from coolchic_b200.enc.training import train as T

acc = {}
def wrap(obj, name, key):
    f = getattr(obj, name)
    def g(*a, **k):
        torch.cuda.synchronize(); t = time.perf_counter()
        r = f(*a, **k)
        torch.cuda.synchronize(); acc[key] = acc.get(key, 0) + time.perf_counter() - t
        return r
    setattr(obj, name, g)
wrap(T, "test", "test")
wrap(T._FusedTrainer, "__init__", "init")
wrap(T._FusedTrainer, "validate", "validate")
wrap(T._FusedTrainer, "pull", "pull")
wrap(T._FusedTrainer, "snapshot", "snapshot")
wrap(T._FusedTrainer, "restore", "restore")
dev = torch.device("cuda:0")
for r in range(4):
    acc.clear()
    out = bench.run_c3x_encode(dev, 0)
    print(round(out["seconds"], 3), [p.get("seconds", p.get("quantize_model_seconds", p.get("warmup_seconds"))) for p in out["phases"]],
          {k: round(v, 3) for k, v in acc.items()}, flush=True)
