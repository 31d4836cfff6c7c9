"""Single-start configs (c1 TSP20 B=512 greedy, c2 TSP50 B=4096 sampling, c3 CVRP100 B=8192 greedy): rollout kernel
time by CUDA events for the shared-memory-resident kernel (default) and the round-1 streaming kernel (impl=STREAM).
usage (GPU box): python profiles/s1_timing.py [c1|c2|c3 ...] [--impl auto|stream] [--reps N]"""
import os, sys, statistics, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib

SPECS = {"c1": ("tsp", 20, 512, "greedy"), "c2": ("tsp", 50, 4096, "sampling"), "c3": ("cvrp", 100, 8192, "greedy")}
ALGO = {"c1": 31838, "c2": 77978, "c3": 156494}


def main():
    which = [a for a in sys.argv[1:] if a in SPECS] or list(SPECS)
    impls = {"auto": _lib.IMPL_AUTO, "stream": _lib.IMPL_STREAM, "coop": _lib.IMPL_COOP}
    if "--impl" in sys.argv:
        k = sys.argv[sys.argv.index("--impl") + 1]
        impls = {k: impls[k]}
    reps = int(sys.argv[sys.argv.index("--reps") + 1]) if "--reps" in sys.argv else 10
    dev = torch.device("cuda", 0)
    for name in which:
        env_name, N, B, dt = SPECS[name]
        torch.manual_seed(1234)
        inst = (P.TSPGenerator if env_name == "tsp" else P.CVRPGenerator)(num_loc=N)(B)
        torch.manual_seed(4321)
        pol = P.AttentionModelPolicy(env_name=env_name).eval().to(dev)
        env = P.get_env(env_name, generator_params=dict(num_loc=N), device=dev)
        td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
        with torch.no_grad():
            h, _ = pol.encoder(td)
        for iname, impl in impls.items():
            events = []
            cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False, _events=events,
                       impl=impl)
            for it in range(2 + reps):
                cfg["noise_seed"] = it
                out = pol._forward_fused(td, env, h, dt, None, 10**6, False, cfg)
            torch.cuda.synchronize()
            ms = statistics.median(a.elapsed_time(b) for a, b in events[2:])
            isteps = out[1].numel()
            print(json.dumps({"config": name, "impl": iname, "kernel_ms": ms, "instance_steps": isteps,
                              "instance_steps_per_s": isteps / ms * 1e3,
                              "algorithmic_GBps": isteps * ALGO[name] / ms / 1e6}), flush=True)


if __name__ == "__main__":
    main()
