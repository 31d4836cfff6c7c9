"""Measured parity per BASELINE.json config -> gpurun_out/parity.json (summarised in profiles/r02_parity.md).

For every config shape: exact-row rate of the CUDA rollout against the oracle (free-running, same seeds / counter
noise), the worst top-2 gap of the oracle at a first divergence (the "tie" the bar admits), and the maximum relative
log-likelihood / reward error of the teacher-forced re-evaluation of EVERY row.  The oracle is the checker here
(test infrastructure), exactly as in tests/.  usage (GPU box): python profiles/parity_report.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import eam_rl4co_b200 as P  # noqa: E402
from eam_rl4co_b200 import _lib  # noqa: E402
from oracle import rollout_oracle as O  # noqa: E402
from helpers import compare_rollout, decoder_weights, inst_to_env_td  # noqa: E402

CASES = [  # name, env, pomo, N, B (oracle-sized), decode types
    ("c1 TSP20 AM", "tsp", False, 20, 512, ["greedy", "sampling"]),
    ("c2 TSP50 AM", "tsp", False, 50, 256, ["greedy", "sampling"]),
    ("c3 CVRP100 AM", "cvrp", False, 100, 96, ["greedy", "sampling"]),
    ("c4 TSP100 POMO", "tsp", True, 100, 12, ["multistart_greedy", "multistart_sampling"]),
    ("c5 CVRP100 POMO", "cvrp", True, 100, 8, ["multistart_greedy", "multistart_sampling"]),
    ("TSP20 POMO", "tsp", True, 20, 64, ["multistart_greedy", "multistart_sampling"]),
    ("CVRP20 POMO", "cvrp", True, 20, 64, ["multistart_greedy", "multistart_sampling"]),
]


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    res = []
    for name, env_name, pomo, N, B, dts in CASES:
        torch.manual_seed(1234)
        inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
        torch.manual_seed(4321)
        pol = (P.POMOPolicy(env_name) if pomo else P.AttentionModelPolicy(env_name=env_name)).eval()
        w = decoder_weights(pol)
        pol = pol.to("cuda")
        env, td = inst_to_env_td(P, env_name, inst, "cuda")
        with torch.no_grad():
            h, _ = pol.encoder(td)
        for dt in dts:
            impls = ([("tc resident", _lib.IMPL_TC, _lib.TC_LAYOUT_RESIDENT), ("tc streaming", _lib.IMPL_TC, _lib.TC_LAYOUT_STREAM),
                      ("fp32", _lib.IMPL_FP32, None)] if pomo else
                     [("coop" if B < 148 * 8 else "stream", _lib.IMPL_AUTO, None)])
            for iname, impl, lay in impls:
                pol.decoder.tc_layout = lay
                out = pol(td.clone(), env, decode_type=dt, noise_seed=2026, impl=impl)
                st = compare_rollout(env_name, inst, h.cpu(), w, not pomo, out, dt, noise_seed=2026, min_exact=0.0,
                                     tie_gap=float("inf"), ll_rtol=1.0)
                rec = dict(config=name, decode_type=dt, impl=iname, **st)
                print(json.dumps(rec), flush=True)
                res.append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "parity.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
