"""TEST INFRASTRUCTURE ONLY — generate tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
Each fixture holds one (env, policy-variant, size) case: instance data, the reference's
decoder weights, its encoder output `h`, and for every decode type the reference's own
actions / reward / log_likelihood / per-step logprobs (+ the first decoder steps' raw
logits and masks).  Seeds follow SURVEY.md §8d: instances 1234, weights 4321, noise 2026.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_runner as RR  # noqa: E402
from oracle import rollout_oracle as O  # noqa: E402
from oracle import synth_weights as SW  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
NOISE_SEED = 2026
CASES = [  # (env, pomo, N, B)
    ("tsp", False, 20, 8), ("tsp", True, 20, 4), ("cvrp", False, 20, 8), ("cvrp", True, 20, 4),
    ("tsp", False, 50, 4), ("cvrp", False, 50, 4), ("tsp", True, 100, 2), ("cvrp", True, 100, 2),
    # BASELINE config c3's own shape (CVRP100 AttentionModel, S = 1, N' = 101: four mask words, the last with five
    # bits) and its TSP twin; single-start decode types only
    ("tsp", False, 100, 4), ("cvrp", False, 100, 4),
]
# encoder fixtures: (env, pomo, N, B) -> reference encoder output for oracle/synth_weights.py weights
ENC_CASES = [("tsp", False, 50, 4), ("cvrp", False, 100, 3), ("tsp", True, 100, 3), ("cvrp", True, 50, 4)]
WEIGHT_SEED = 97
DECODE_TYPES = ["greedy", "sampling", "multistart_greedy", "multistart_sampling"]
TRACE_STEPS = 3


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    for env_name, pomo, N, B in CASES:
        torch.manual_seed(1234)
        inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
        pol = RR.make_policy(env_name, pomo, seed=4321)
        w = RR.decoder_weights(pol)
        rec = {f"inst.{k}": v.numpy() for k, v in inst.items()}
        rec.update({f"w.{k}": v.numpy() for k, v in w.items()})
        rec["meta.use_graph_context"] = np.array(not pomo)
        rec["meta.noise_seed"] = np.array(NOISE_SEED)
        for dt in DECODE_TYPES:
            if N >= 100 and pomo != dt.startswith("multistart"):
                continue        # N = 100: POMO fixtures hold the multistart types, AM fixtures the single-start ones
            trace = {}
            out = RR.run_reference(env_name, inst, pol, dt, noise_seed=NOISE_SEED, trace=trace)
            if "h" not in rec:
                rec["h"] = out["hidden"].node_embeddings.numpy()
            rec[f"{dt}.actions"] = out["actions"].numpy().astype(np.int16)
            rec[f"{dt}.reward"] = out["reward"].numpy()
            rec[f"{dt}.log_likelihood"] = out["log_likelihood"].numpy()
            for t in range(min(TRACE_STEPS, len(trace["logits"]))):
                if N < 100:
                    rec[f"{dt}.logits.{t}"] = trace["logits"][t].numpy()
                    rec[f"{dt}.mask.{t}"] = trace["mask"][t].numpy()
            # teacher-forced evaluate of the same actions (decoding.py:452-465)
            S = N if dt.startswith("multistart") else None
            ev = RR.run_reference(env_name, inst, pol, "evaluate", actions=out["actions"], num_starts=S)
            rec[f"{dt}.eval_log_likelihood"] = ev["log_likelihood"].numpy()
        name = f"{env_name}{N}_{'pomo' if pomo else 'am'}_B{B}.npz"
        np.savez_compressed(os.path.join(OUT, name), **rec)
        print(name, os.path.getsize(os.path.join(OUT, name)) // 1024, "KiB")


def encoder_fixtures():
    """h = reference encoder (am/encoder.py:70-91, attnnet.py:16-103, ops.py:32-56, env_embeddings/init.py) on seeded
    instances with synthetic weights; plus the greedy tour of the full reference policy on top of it."""
    for env_name, pomo, N, B in ENC_CASES:
        torch.manual_seed(4242)
        inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
        pol = RR.make_policy(env_name, pomo, seed=4321)
        sd = SW.synth_state_dict(pol.state_dict(), WEIGHT_SEED)
        pol.load_state_dict(sd, strict=True)
        pol.eval()
        dt = "multistart_greedy" if pomo else "greedy"
        out = RR.run_reference(env_name, inst, pol, dt)
        rec = {f"inst.{k}": v.numpy() for k, v in inst.items()}
        rec["h"] = out["hidden"].node_embeddings.numpy()
        rec["meta.weight_seed"] = np.array(WEIGHT_SEED)
        rec["meta.signature"] = np.array(SW.signature(pol.state_dict()))
        rec["meta.pomo"] = np.array(pomo)
        rec[f"{dt}.actions"] = out["actions"].numpy().astype(np.int16)
        rec[f"{dt}.reward"] = out["reward"].numpy()
        rec[f"{dt}.log_likelihood"] = out["log_likelihood"].numpy()
        name = f"enc_{env_name}{N}_{'pomo' if pomo else 'am'}_B{B}.npz"
        np.savez_compressed(os.path.join(OUT, name), **rec)
        print(name, os.path.getsize(os.path.join(OUT, name)) // 1024, "KiB")


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "rollout"):
        main()
    if which in ("all", "encoder"):
        encoder_fixtures()
