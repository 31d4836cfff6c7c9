"""ll_gap (rollout log-likelihood vs differentiable re-evaluation) and loss over consecutive c5 training steps, for the
combinations graphed encoder x fused Adam.  usage (GPU box): python profiles/debug/ll_gap.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import train

dev = torch.device("cuda", 0)
for graphed in (False, True):
    for fused in (False, True):
        torch.manual_seed(4321)
        pol = P.POMOPolicy("cvrp").to(dev).train()
        opt = torch.optim.Adam(pol.parameters(), lr=1e-4, fused=fused)
        env = P.get_env("cvrp", generator_params=dict(num_loc=100), device=dev)
        gen = P.DeviceGenerator("cvrp", num_loc=100, seed=7, device=dev)
        enc = train.graphed_encoder(pol, env.reset(gen(64, instance_offset=0))) if graphed else None
        gaps = []
        for it in range(14):
            td = env.reset(gen(64, instance_offset=it * 64))
            st = train.pomo_train_step(pol, env, td, opt, noise_seed=it, instance_offset=it * 64, encoder=enc)
            gaps.append((round(st["ll_gap"], 5), round(st["loss"], 3), round(st["reward"], 2)))
        print("graphed", graphed, "fused adam", fused, gaps[:3], "...", gaps[-3:])
