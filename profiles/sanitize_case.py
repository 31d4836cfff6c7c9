"""Tiny end-to-end exercise of every product kernel for compute-sanitizer (profiles/r02_sanitize.sh):
TSP20 / CVRP20 multistart rollouts on the tcgen05 kernel (greedy, sampling, evaluate), the fp32 multistart kernel,
single-start rollouts on the cooperative and streaming kernels, the step-granular kernels, the encoder kernels,
best-of-starts, the state replay and one training step (query + evaluate kernels, forward and backward)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import _lib, train

dev = "cuda"
for env_name in ("tsp", "cvrp"):
    N, B = 20, 3
    torch.manual_seed(1)
    inst = (P.TSPGenerator if env_name == "tsp" else P.CVRPGenerator)(num_loc=N)(B)
    torch.manual_seed(2)
    pomo = P.POMOPolicy(env_name).eval().to(dev)
    am = P.AttentionModelPolicy(env_name=env_name).eval().to(dev)
    env = P.get_env(env_name, generator_params=dict(num_loc=N), device=dev)
    td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
    for dt in ("multistart_greedy", "multistart_sampling"):
        for impl in (_lib.IMPL_TC, _lib.IMPL_FP32):
            out = pomo(td.clone(), env, decode_type=dt, noise_seed=3, impl=impl)
    ev = pomo(td.clone(), env, actions=out["actions"], num_starts=N)
    P.best_of_starts(dict(out), 1, N)
    for dt in ("greedy", "sampling"):
        for impl in (_lib.IMPL_COOP, _lib.IMPL_STREAM):
            o1 = am(td.clone(), env, decode_type=dt, noise_seed=4, impl=impl)
    o2 = am(td.clone(), env, decode_type="greedy", fused=False)          # step-granular kernels
    o3 = pomo(td.clone(), env, decode_type="multistart_greedy", fused=False)
    st = train.teacher_forced_states(env, td, out["actions"], N)
    # training step: query rows, evaluate forward / backward (all tcgen05 kernels), query gradient
    pomo.train()
    stats = train.pomo_train_step(pomo, env, td.clone(), optimizer=None, noise_seed=5)
    pomo.eval()
    torch.cuda.synchronize()
    print(env_name, "ok", out["reward"].mean().item(), o1["reward"].mean().item(), o2["reward"].mean().item(), stats["loss"])
print("sanitize case done")
