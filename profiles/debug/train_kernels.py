"""c5 training step: per-kernel device time of one `pomo_train_step` from torch.profiler (CUPTI), to see what the
PyTorch glue around the evaluate kernels costs.  usage (GPU box): python profiles/debug/train_kernels.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import profile, ProfilerActivity
import eam_rl4co_b200 as P
from eam_rl4co_b200 import train as T


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    dev = torch.device("cuda", 0)
    torch.manual_seed(4321)
    pol = P.POMOPolicy("cvrp").to(dev).train()
    opt = torch.optim.Adam(pol.parameters(), lr=1e-4, fused=True)
    env = P.get_env("cvrp", generator_params=dict(num_loc=100), device=dev)
    inst = {k: v for k, v in P.CVRPGenerator(num_loc=100)(B).items() if k != "capacity"}
    td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
    encode = T.graphed_encoder(pol, td) if os.environ.get("R4_GRAPH_ENCODER", "1") == "1" else None
    for it in range(3):
        T.pomo_train_step(pol, env, td.clone(), opt, noise_seed=it, encoder=encode)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for it in range(3):
            T.pomo_train_step(pol, env, td.clone(), opt, noise_seed=10 + it, encoder=encode)
        torch.cuda.synchronize()
    ev = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type == torch.autograd.DeviceType.CUDA]
    ev.sort(key=lambda e: -e.device_time_total)
    tot = sum(e.device_time_total for e in ev)
    print(f"device time per step {tot / 3e3:.2f} ms over {sum(e.count for e in ev) // 3} launches")
    for e in ev[:45]:
        print(f"{e.device_time_total / 3e3:8.3f} ms  x{e.count // 3:<4d} {e.key[:150]}")


if __name__ == "__main__":
    main()
