"""CPU, build container only: the oracle against the LIVE reference (/root/reference run
unmodified through oracle/ref_loader.py).  Skipped where the reference is absent."""
import pytest
import torch

from oracle.ref_loader import reference_available
from oracle import rollout_oracle as O

pytestmark = pytest.mark.skipif(not reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("env_name", ["tsp", "cvrp"])
@pytest.mark.parametrize("pomo", [False, True])
def test_rollout_bit_identical(env_name, pomo):
    from oracle import ref_runner as RR
    torch.set_num_threads(4)
    torch.manual_seed(7)
    N, B = 15, 6
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    pol = RR.make_policy(env_name, pomo)
    w = RR.decoder_weights(pol)
    for dt in ("greedy", "sampling", "multistart_greedy", "multistart_sampling"):
        ref = RR.run_reference(env_name, inst, pol, dt, noise_seed=11)
        h = ref["hidden"].node_embeddings
        with torch.inference_mode():
            mine = O.rollout(env_name, inst, h, w, dt, use_graph_context=not pomo, noise_seed=11)
        assert torch.equal(ref["actions"], mine["actions"])
        assert torch.equal(ref["reward"], mine["reward"])
        assert torch.equal(ref["log_likelihood"], mine["log_likelihood"])


def test_env_steps_bit_identical():
    """Random feasible action sequences through the reference env vs the oracle env functions."""
    from oracle import ref_runner as RR
    torch.manual_seed(3)
    for env_name in ("tsp", "cvrp"):
        N, B = 12, 16
        inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
        env = RR.make_env(env_name, N)
        td = env.reset(RR.inst_to_td(env_name, inst))
        mine = O.tsp_reset(inst["locs"]) if env_name == "tsp" else O.cvrp_reset(
            inst["locs"], inst["depot"], inst["demand"])
        step = O.tsp_step if env_name == "tsp" else O.cvrp_step
        while not td["done"].all():
            a = torch.multinomial(td["action_mask"].float(), 1).squeeze(1)
            td.set("action", a)
            td = env.step(td)["next"]
            mine["action"] = a
            mine = step(mine)
            for k in ("action_mask", "current_node", "done"):
                assert torch.equal(td[k], mine[k]), k
            if env_name == "cvrp":
                assert torch.equal(td["used_capacity"], mine["used_capacity"])
                assert torch.equal(td["visited"], mine["visited"])
