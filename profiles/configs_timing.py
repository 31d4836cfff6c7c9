"""Rollout throughput of every BASELINE.json config (c1-c5 rollout part) on one B200: the fused rollout kernel alone
(CUDA events around rl4co_rollout), device-resident, against SURVEY.md 8(d)'s algorithmic bytes per instance-step and
MEASURED_PEAKS.json's HBM figure.  bench.py reports c4 only (the headline); this is the side table DESIGN.md quotes.
usage (GPU box): python profiles/configs_timing.py"""
import json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
import bench

CONFIGS = [  # name, env, N, B, policy, decode_type, algorithmic bytes per instance-step
    ("c1 TSP20 AM greedy B=512", "tsp", 20, 512, "am", "greedy", 31838),
    ("c2 TSP50 AM sampling B=4096", "tsp", 50, 4096, "am", "sampling", 77978),
    ("c3 CVRP100 AM greedy B=8192", "cvrp", 100, 8192, "am", "greedy", 156494),
    ("c4 TSP100 POMO greedy S=100 B=1024", "tsp", 100, 1024, "pomo", "multistart_greedy", 2302),
    ("c5 CVRP100 POMO sampling S=100 B=64 (train rollout)", "cvrp", 100, 64, "pomo", "multistart_sampling", 2513),
    ("c5' CVRP100 POMO sampling S=100 B=1024", "cvrp", 100, 1024, "pomo", "multistart_sampling", 2513),
]


def main():
    dev = torch.device("cuda", 0)
    peak, src = bench.peaks()
    out = []
    for name, env_name, N, B, pk, dt, abytes in CONFIGS:
        torch.manual_seed(1234)
        gen = P.TSPGenerator(num_loc=N) if env_name == "tsp" else P.CVRPGenerator(num_loc=N)
        inst = {k: v for k, v in gen(B).items() if k != "capacity"}
        torch.manual_seed(4321)
        pol = (P.POMOPolicy(env_name) if pk == "pomo" else P.AttentionModelPolicy(env_name=env_name)).eval().to(dev)
        env = P.get_env(env_name, generator_params=dict(num_loc=N), device=dev)
        td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
        with torch.no_grad():
            h, _ = pol.encoder(td)
        events = []
        cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False, noise_seed=2026,
                   _events=events)
        isteps = None
        for it in range(3 + 10):
            if it == 3:
                torch.cuda.synchronize()
                events.clear()
            _, acts, _, _, _ = pol._forward_fused(td.clone(), env, h, dt, None, 10**6, False, cfg)
            isteps = acts.shape[0] * acts.shape[1]
        torch.cuda.synchronize()
        ms = statistics.median(a.elapsed_time(b) for a, b in events)
        rate = isteps / (ms * 1e-3)
        gbs = rate * abytes / 1e9
        row = {"config": name, "rows": acts.shape[0], "T": acts.shape[1], "kernel_ms": ms,
               "instance_steps_per_s": rate, "achieved_GBs": gbs, "peak_GBs": peak, "frac": gbs / peak}
        out.append(row)
        print(json.dumps(row), flush=True)
    print(f"peak source: {src}")


if __name__ == "__main__":
    main()
