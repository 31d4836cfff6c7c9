"""In-kernel phase timing of the tensor-path rollout kernel (debug counter rl4co_rollout_args.phase_cycles):
clock64 cycles thread 0 of CTA 0 spends in each phase, accumulated over its instances and steps.
usage: profiles/phase.sh [B] [tsp|cvrp] [decode_type]   (needs the -DR4_PHASE_PROF build; the product build reports zeros)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
import bench

NAMES = ["Q(0) build+team sync", "wait scores(0)", "softmax(0) +team vote", "issue PV(0)+Q(1) build", "wait PV(0)",
         "team sync+issue scores(1)+drain O(0)", "wait scores(1)", "softmax(1) +team vote", "issue PV(1)", "wait PV(1)",
         "drain O(1)+CTA sync+heads->TMEM+sync", "logits issue+wait", "select", "env step+sync", "instance prologue", "-"]

def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    env_name = sys.argv[2] if len(sys.argv) > 2 else "tsp"
    dt = sys.argv[3] if len(sys.argv) > 3 else "multistart_greedy"
    dev = torch.device("cuda", 0)
    if env_name == "tsp":
        pol = bench.make_policy().to(dev)
        env = P.TSPEnv(generator_params=dict(num_loc=bench.N_NODES), device=dev)
        inst = {"locs": bench.make_instances(B)}
    else:
        torch.manual_seed(4321)
        pol = P.POMOPolicy("cvrp").eval().to(dev)
        env = P.get_env("cvrp", generator_params=dict(num_loc=bench.N_NODES), device=dev)
        torch.manual_seed(1234)
        inst = {k: v for k, v in P.CVRPGenerator(num_loc=bench.N_NODES)(B).items() if k != "capacity"}
    td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
    with torch.no_grad():
        h, _ = pol.encoder(td)
    for it in range(3):
        pc = torch.zeros(16 + 64, dtype=torch.int64, device=dev)
        cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False, _phase_cycles=pc,
                   noise_seed=2026)
        try:
            _, acts, _, _, _ = pol._forward_fused(td.clone(), env, h, dt, None, 10**6, False, cfg)
        except AssertionError as e:     # experiment builds with broken numerics still time correctly
            print("assertion:", str(e)[:80])
        torch.cuda.synchronize()
    c = pc.tolist()
    n_inst = (B + 147) // 148
    steps = n_inst * (acts.shape[1] - 1)      # CVRP: upper bound (instances of CTA 0 may finish before the batch maximum)
    misses, c[15] = c[15], 0
    tot = sum(c[:16])
    print(f"CTA 0: {n_inst} instances, {steps} decode steps, {tot} cycles total, {tot / steps:.0f} cycles/step")
    for name, v in zip(NAMES, c):
        print(f"  {name:34s} {v / steps:9.0f} cyc/step  {100.0 * v / tot:5.1f}%")
    print(f"  exact-form fallbacks of team 0 (missed softmax reference): {misses} in {2 * steps} head passes")
    if any(c[16:]):     # -DR4_WARP_PROF builds: per-warp select / env step / wait at the step-end barrier
        for w in range(16):
            sel, envs, wait = (c[16 + 4 * w + k] / steps for k in range(3))
            print(f"  warp {w:2d} (q={w % 4}, j={w // 4}): select {sel:6.0f}  env {envs:6.0f}  step-end wait {wait:6.0f}")

if __name__ == "__main__":
    main()
