# one launch of the fixed-point ARM kernel at 768x512 for ncu
import sys
sys.path.insert(0, ".")
import torch, bench
print(bench.time_arm_int(torch.device("cuda:0"), True))
