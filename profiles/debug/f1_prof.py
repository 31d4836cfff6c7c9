"""torch.profiler kernel table of one c5 training step (which kernels the re-evaluation spends its time in)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import train as T
from torch.profiler import profile, ProfilerActivity

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
torch.manual_seed(4321)
pol = P.POMOPolicy("cvrp").to(dev).train()
env = P.get_env("cvrp", generator_params=dict(num_loc=100), device=dev)
torch.manual_seed(1234)
inst = {k: v for k, v in P.CVRPGenerator(num_loc=100)(B).items() if k != "capacity"}
td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
for it in range(3):
    T.pomo_train_step(pol, env, td.clone(), optimizer=None, noise_seed=it)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    T.pomo_train_step(pol, env, td.clone(), optimizer=None, noise_seed=11)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=200, max_name_column_width=70))
