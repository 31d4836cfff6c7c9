"""c5: one POMO/REINFORCE training step on CVRP100 (B instances x 100 starts, reference default B = 64 per device):
CUDA-event split into the sampled rollout (fused kernel), the differentiable teacher-forced re-evaluation (PyTorch,
forward), backward, and the optimiser step.  usage (GPU box): python profiles/train_timing.py [B]"""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
from eam_rl4co_b200 import train as T


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    dev = torch.device("cuda", 0)
    torch.manual_seed(4321)
    pol = P.POMOPolicy("cvrp").to(dev).train()
    opt = torch.optim.Adam(pol.parameters(), lr=1e-4, fused=True)
    env = P.get_env("cvrp", generator_params=dict(num_loc=100), device=dev)
    torch.manual_seed(1234)
    inst = {k: v for k, v in P.CVRPGenerator(num_loc=100)(B).items() if k != "capacity"}
    td = env.reset(P.TensorDict({k: v.to(dev) for k, v in inst.items()}, batch_size=[B]))
    ev = lambda: torch.cuda.Event(enable_timing=True)
    encode = T.graphed_encoder(pol, td) if os.environ.get("R4_GRAPH_ENCODER", "1") == "1" else None
    rows = []
    for it in range(3 + 5):
        e = [ev() for _ in range(6)]
        e[0].record()
        with torch.no_grad():
            out = pol(td.clone(), env, decode_type="multistart_sampling", noise_seed=it)
        e[1].record()
        S = out["actions"].shape[0] // B
        states = T.teacher_forced_states(env, td, out["actions"], S)
        e[2].record()
        ll = T.evaluate_log_likelihood(pol, td, env, out["actions"], S, states=states, hidden=encode(td) if encode else None)
        loss, _ = T.reinforce_loss(out["reward"], ll, S)
        e[3].record()
        opt.zero_grad(set_to_none=True)
        loss.backward()
        e[4].record()
        torch.nn.utils.clip_grad_norm_(pol.parameters(), 1.0)
        opt.step()
        e[5].record()
        torch.cuda.synchronize()
        if it >= 3:
            rows.append([e[i].elapsed_time(e[i + 1]) for i in range(5)] + [out["actions"].numel()])
    med = [statistics.median(r[i] for r in rows) for i in range(5)]
    isteps = rows[-1][5]
    tot = sum(med)
    print(f"CVRP100 POMO train step, B={B} x S={S}, T={out['actions'].shape[1]}: total {tot:.1f} ms "
          f"({isteps / tot * 1e3:.3e} instance-steps/s)")
    names = ["policy(): encoder + sampled rollout (fused kernel)", "state replay (rl4co_replay_states, one launch)",
             "teacher-forced re-evaluation, forward (rl4co_eval_ll_fwd + torch glue)", "backward (rl4co_eval_ll_bwd + torch autograd glue)", "clip + Adam step"]
    for name, v in zip(names, med):
        print(f"  {name:52s} {v:8.2f} ms  {100 * v / tot:5.1f}%")
    print(f"  peak memory {torch.cuda.max_memory_allocated() / 2**30:.2f} GiB")


if __name__ == "__main__":
    main()
