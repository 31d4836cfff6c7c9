import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, bench
import eam_rl4co_b200 as P
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
pol = bench.make_policy().to(dev)
env = P.TSPEnv(generator_params=dict(num_loc=100), device=dev)
td = env.reset(P.TensorDict({"locs": bench.make_instances(B).to(dev)}, batch_size=[B]))
for _ in range(2): pol(td.clone(), env, decode_type="multistart_greedy")
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    pol(td.clone(), env, decode_type="multistart_greedy"); torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:6]:
    print(f"{e.device_time_total/1e3:9.3f} ms  x{e.count:3d}  {e.key[:100]}")
