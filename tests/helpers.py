"""Shared helpers for the parity tests (oracle = checker, never the product path)."""
import glob
import os

import numpy as np
import torch

from oracle import rollout_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# A first divergence counts as a near-tie when the oracle's own gap between the two candidates is below
# 1e-5 * max(1, |logit|); clipped logits are bounded by tanh_clipping = 10, hence 1e-4.  The gaps
# actually admitted per config are recorded in profiles/r02_parity.md (worst seen: see there).
TIE_GAP = 1e-4
LL_RTOL = 1e-5      # north_star: rewards and log-likelihoods within 1e-5 relative (fp32)


def golden_files():
    """rollout fixtures (oracle/make_golden.py main)"""
    return sorted(p for p in glob.glob(os.path.join(GOLDEN, "*.npz")) if not os.path.basename(p).startswith("enc_"))


def encoder_golden_files():
    """encoder fixtures (oracle/make_golden.py encoder_fixtures)"""
    return sorted(glob.glob(os.path.join(GOLDEN, "enc_*.npz")))


def load_encoder_golden(path, P, device="cpu"):
    """-> dict(env, inst, h, policy) with the mirror policy carrying the fixture's synthetic weights
    (oracle/synth_weights.py, loaded strict=True; the key/shape signature must equal the reference's)."""
    from oracle import synth_weights as SW
    z = np.load(path)
    name = os.path.basename(path)
    env = "tsp" if name.startswith("enc_tsp") else "cvrp"
    pomo = bool(z["meta.pomo"])
    inst = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("inst.")}
    pol = (P.POMOPolicy(env) if pomo else P.AttentionModelPolicy(env_name=env)).eval()
    assert SW.signature(pol.state_dict()) == str(z["meta.signature"]), "state_dict keys/shapes differ from the reference"
    pol.load_state_dict(SW.synth_state_dict(pol.state_dict(), int(z["meta.weight_seed"])), strict=True)
    pol = pol.eval().to(device)
    return dict(env=env, pomo=pomo, inst=inst, h=torch.from_numpy(z["h"]), policy=pol, z=z, name=name)


def encoder_input(inst, device="cpu"):
    """what the init embedding reads after env.reset (tsp/env.py:90-115, cvrp/env.py:102-130), without the env"""
    if "depot" in inst:
        return {"locs": torch.cat((inst["depot"][:, None, :], inst["locs"]), 1).to(device),
                "demand": inst["demand"].to(device)}
    return {"locs": inst["locs"].to(device)}


def load_golden(path):
    z = np.load(path)
    name = os.path.basename(path)
    env = "tsp" if name.startswith("tsp") else "cvrp"
    inst = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("inst.")}
    w = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w.")}
    return dict(env=env, inst=inst, w=w, h=torch.from_numpy(z["h"]),
                use_graph_context=bool(z["meta.use_graph_context"]),
                noise_seed=int(z["meta.noise_seed"]), z=z, name=name)


def decoder_weights(policy):
    return {k: v.detach().cpu().clone() for k, v in policy.decoder.state_dict().items()}


def inst_to_env_td(P, env_name, inst, device):
    """generator-style instance dict -> env.reset(td) on `device`."""
    env = P.get_env(env_name, generator_params=dict(num_loc=inst["locs"].shape[1]), device=device)
    td = P.TensorDict({k: v.to(device) for k, v in inst.items()}, batch_size=[inst["locs"].shape[0]])
    return env, env.reset(td)


def rel_close(a, b, rtol):
    a, b = a.double().cpu(), b.double().cpu()
    return bool(((a - b).abs() <= rtol * b.abs().clamp_min(1.0)).all())


def compare_rollout(env_name, inst, h, w, use_gc, got, decode_type, noise_seed=0, num_starts=None,
                    instance_offset=0, min_exact=0.97, tie_gap=TIE_GAP, ll_rtol=LL_RTOL):
    """Tie-aware parity of a CUDA rollout `got` (dict with actions, reward, log_likelihood) against
    the oracle.  Returns a stats dict; asserts on any non-tie divergence."""
    with torch.inference_mode():
        ref = O.rollout(env_name, inst, h, w, decode_type, num_starts=num_starts,
                        use_graph_context=use_gc, noise_seed=noise_seed, instance_offset=instance_offset)
    a_got = got["actions"].cpu()
    a_ref = ref["actions"]
    T = max(a_got.shape[1], a_ref.shape[1])
    pad = lambda a: torch.nn.functional.pad(a, (0, T - a.shape[1]))
    a_got_p, a_ref_p = pad(a_got), pad(a_ref)
    same_rows = (a_got_p == a_ref_p).all(1)
    exact = same_rows.float().mean().item()
    # teacher-force the oracle on the CUDA actions: checks numerics of EVERY row, diverged or not
    S = ref["num_starts"]
    with torch.inference_mode():
        ev = O.rollout(env_name, inst, h, w, actions=a_got, num_starts=S if S else None,
                       use_graph_context=use_gc)
    rel = lambda a, b: ((a.double().cpu() - b.double()).abs() / b.double().abs().clamp_min(1.0)).max().item()
    ll_err, r_err = rel(got["log_likelihood"], ev["log_likelihood"]), rel(got["reward"], ev["reward"])
    assert ll_err <= ll_rtol, ("log-likelihood off", ll_err)
    assert r_err <= ll_rtol, ("reward off", r_err)
    # every diverged row must diverge at a near-tie of the oracle's own scores
    worst_gap = 0.0
    for r in torch.nonzero(~same_rows).flatten().tolist():
        t = int(torch.nonzero(a_got_p[r] != a_ref_p[r])[0])
        lp_ref_choice = ref["logprobs"][r, t].item()
        lp_got_choice = ev["logprobs"][r, t].item() if t < ev["logprobs"].shape[1] else float("-inf")
        if "sampling" in decode_type:
            n = a_ref.new_tensor(0)  # gumbel perturbation shifts both candidates; compare perturbed
            B = inst["locs"].shape[0]
            g = O.noise_gumbel(noise_seed, torch.tensor([r % B + instance_offset]), torch.tensor([r // B]),
                               t, h.shape[1])[0]
            gap = abs((lp_ref_choice + g[a_ref_p[r, t]].item()) - (lp_got_choice + g[a_got_p[r, t]].item()))
        else:
            gap = abs(lp_ref_choice - lp_got_choice)
        worst_gap = max(worst_gap, gap)
        assert gap < tie_gap, f"row {r} diverges at step {t} with gap {gap} (not a tie)"
    assert exact >= min_exact, f"exact-match rate {exact} < {min_exact}"
    return dict(exact=exact, worst_gap=worst_gap, T_got=a_got.shape[1], T_ref=a_ref.shape[1], ll_rel=ll_err,
                reward_rel=r_err, rows=int(a_got.shape[0]), diverged=int((~same_rows).sum()))
