import os, sys
sys.path.insert(0, "/root/repo")
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import train as T
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
B = 64
torch.manual_seed(4321)
pol = P.POMOPolicy("cvrp").to(dev).train()
env = P.get_env("cvrp", generator_params=dict(num_loc=100), device=dev)
torch.manual_seed(1234)
inst = {k: v for k, v in P.CVRPGenerator(num_loc=100)(B).items() if k != "capacity"}
td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
def step(it):
    with torch.no_grad():
        out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=it)
    S = out["actions"].shape[0] // B
    ll = T.evaluate_log_likelihood(pol, td, env, out["actions"], S)
    loss, _ = T.reinforce_loss(out["reward"], ll, S)
    loss.backward()
for it in range(3): step(it)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(5); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=70))
