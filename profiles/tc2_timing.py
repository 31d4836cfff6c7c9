"""Resident (rollout_tc_kernel) vs streaming (rollout_tc2_kernel) tensor-path rollout, TSP100 POMO greedy, kernel time by CUDA
events.  usage (GPU box): python profiles/tc2_timing.py [B ...]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib

def main():
    Bs = [int(a) for a in sys.argv[1:] if a.isdigit()] or [148, 296, 1024]
    env_name = "cvrp" if "cvrp" in sys.argv else "tsp"
    dt = "multistart_sampling" if "sampling" in sys.argv else "multistart_greedy"
    dev = torch.device("cuda", 0)
    torch.manual_seed(4321)
    pol = P.POMOPolicy(env_name).eval().to(dev)
    env = P.get_env(env_name, generator_params=dict(num_loc=100), device=dev)
    for B in Bs:
        torch.manual_seed(1234)
        inst = (P.TSPGenerator if env_name == "tsp" else P.CVRPGenerator)(num_loc=100)(B)
        td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
        with torch.no_grad():
            h, _ = pol.encoder(td)
        for name, lay in (("resident", _lib.TC_LAYOUT_RESIDENT), ("stream", _lib.TC_LAYOUT_STREAM)):
            pol.decoder.tc_layout = lay
            events = []
            cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False, _events=events)
            for it in range(3 + 8):
                cfg["noise_seed"] = it
                out = pol._forward_fused(td, env, h, dt, None, 10**6, False, cfg)
            torch.cuda.synchronize()
            ms = statistics.median(a.elapsed_time(b) for a, b in events[3:])
            isteps = out[1].numel()
            print(f"{env_name} {dt} B={B:5d} {name:9s} kernel {ms:8.3f} ms  {isteps / ms * 1e3:.3e} instance-steps/s", flush=True)

if __name__ == "__main__":
    main()
