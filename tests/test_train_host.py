"""CPU: the differentiable teacher-forced evaluate of eam_rl4co_b200/train.py (pure PyTorch part) against autograd
through the oracle's sequential evaluate-mode rollout (reference: decoding.py:452-465, constructive/base.py:204-229,
reinforce.py:97-106, baselines.py:57-61), and the 2-rank gradient all-reduce of the training step."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import eam_rl4co_b200 as P
from eam_rl4co_b200 import sharding, train
from oracle import rollout_oracle as O


def _oracle_states(env_name, inst, actions, S):
    """Replay actions with the oracle's env functions -> the `states` dict train.batched_log_likelihood takes."""
    td = O.tsp_reset(inst["locs"]) if env_name == "tsp" else O.cvrp_reset(inst["locs"], inst["depot"], inst["demand"])
    step = O.tsp_step if env_name == "tsp" else O.cvrp_step
    td = {k: O.batchify(v, S) for k, v in td.items()}
    td["action"] = actions[:, 0]
    td = step(td)
    cur, rem, mask = [], [], []
    for t in range(1, actions.shape[1]):
        cur.append(td["current_node"].reshape(-1))
        mask.append(td["action_mask"])
        if env_name == "cvrp":
            rem.append((td["vehicle_capacity"] - td["used_capacity"]).reshape(-1))
        td["action"] = actions[:, t]
        td = step(td)
    bs = lambda x: train._to_bs(x, S)
    st = {"first": bs(actions[:, 0]), "cur": bs(torch.stack(cur, 1)), "mask": bs(torch.stack(mask, 1)),
          "taken": bs(actions[:, 1:])}
    if env_name == "cvrp":
        st["rem"] = bs(torch.stack(rem, 1))
    return st


def _setup(env_name, B, N, use_graph_context=False):
    torch.manual_seed(11)
    inst = O.generate_tsp(B, N) if env_name == "tsp" else O.generate_cvrp(B, N)
    torch.manual_seed(12)
    pol = P.POMOPolicy(env_name, use_graph_context=use_graph_context)
    enc_in = dict(inst)
    if env_name == "cvrp":
        enc_in = {"locs": torch.cat((inst["depot"][:, None, :], inst["locs"]), 1), "demand": inst["demand"]}
    return inst, pol, enc_in


@pytest.mark.parametrize("env_name,B,N,gc", [("tsp", 3, 10, False), ("tsp", 2, 20, True), ("cvrp", 3, 10, False),
                                             ("cvrp", 2, 12, True)])
def test_batched_log_likelihood_and_gradients_match_the_oracle(env_name, B, N, gc):
    inst, pol, enc_in = _setup(env_name, B, N, gc)
    S = N
    names = [k for k, _ in pol.decoder.named_parameters()]
    w = {k: v.detach().clone().requires_grad_(True) for k, v in pol.decoder.state_dict().items()}
    with torch.no_grad():
        h0, _ = pol.encoder(enc_in)
        smp = O.rollout(env_name, inst, h0, {k: v.detach() for k, v in w.items()}, "multistart_sampling",
                        use_graph_context=gc, noise_seed=5)
    actions, reward = smp["actions"], smp["reward"]

    # oracle: sequential evaluate with autograd
    h_o = h0.clone().requires_grad_(True)
    ev = O.rollout(env_name, inst, h_o, w, actions=actions, num_starts=S, use_graph_context=gc)
    loss_o, adv_o = O.reinforce_shared_baseline_loss(reward, ev["log_likelihood"], S)
    loss_o.backward()

    # product: one batched pass
    h_p = h0.clone().requires_grad_(True)
    st = _oracle_states(env_name, inst, actions, S)
    logp = train.batched_log_likelihood(pol.decoder, h_p, st, pol.tanh_clipping, pol.temperature)
    ll = logp.sum(-1).transpose(0, 1).reshape(-1)
    loss_p, adv_p = train.reinforce_loss(reward, ll, S)
    loss_p.backward()

    assert torch.allclose(ll, ev["log_likelihood"], rtol=1e-5, atol=1e-5)
    assert torch.allclose(ll, smp["log_likelihood"], rtol=1e-5, atol=1e-5)      # evaluate reproduces the sampling pass
    assert torch.allclose(adv_p, adv_o)
    assert torch.allclose(loss_p, loss_o, rtol=1e-5, atol=1e-7)
    assert torch.allclose(h_p.grad, h_o.grad, rtol=1e-4, atol=1e-6)
    params = dict(pol.decoder.named_parameters())
    checked = 0
    for k in names:
        if w[k].grad is None:
            assert params[k].grad is None or params[k].grad.abs().max() == 0, k
            continue
        assert torch.allclose(params[k].grad, w[k].grad, rtol=1e-4, atol=1e-6), k
        checked += 1
    assert checked >= 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _grads_on_shard(env_name, inst, pol, enc_in, actions, reward, lo, hi, S, group_active):
    """Loss + backward of the training step on instances [lo, hi) (rollout results given), then the step's
    single all-reduce."""
    B = inst["locs"].shape[0]
    sl = lambda x: train._to_bs(x, S)[lo:hi].transpose(0, 1).reshape((hi - lo) * S, *x.shape[1:])
    sub = {k: v[lo:hi] for k, v in inst.items()}
    sub_enc = {k: v[lo:hi] for k, v in enc_in.items()}
    a, r = sl(actions), sl(reward)
    h, _ = pol.encoder(sub_enc)
    st = _oracle_states(env_name, sub, a, S)
    ll = train.batched_log_likelihood(pol.decoder, h, st).sum(-1).transpose(0, 1).reshape(-1)
    loss, _ = train.reinforce_loss(r, ll, S)
    params = [p for p in pol.parameters() if p.requires_grad]
    for p in params:
        p.grad = None
    loss.backward()
    grads = [p.grad if p.grad is not None else torch.zeros_like(p) for p in params]
    extras = sharding.allreduce_mean_(grads, torch.stack([loss.detach(), r.mean()])) if group_active else \
        torch.stack([loss.detach(), r.mean()])
    return grads, extras


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        inst, pol, enc_in, actions, reward = _problem()
        B = inst["locs"].shape[0]
        lo, hi = sharding.shard_bounds(B, rank, world)
        grads, extras = _grads_on_shard("tsp", inst, pol, enc_in, actions, reward, lo, hi, 8, True)
        if rank == 0:
            ret["grads"] = [g.clone() for g in grads]
            ret["extras"] = extras.clone()
    finally:
        dist.destroy_process_group()


def _problem():
    inst, pol, enc_in = _setup("tsp", 4, 8)
    w = {k: v.detach().clone() for k, v in pol.decoder.state_dict().items()}
    with torch.no_grad():
        h0, _ = pol.encoder(enc_in)
        smp = O.rollout("tsp", inst, h0, w, "multistart_sampling", use_graph_context=False, noise_seed=9)
    return inst, pol, enc_in, smp["actions"], smp["reward"]


def test_two_rank_training_gradients_equal_single_rank():
    """DDP mean semantics (trainer.py:86-89): equal shards -> the all-reduced mean of per-shard gradients of the
    per-shard mean loss equals the gradient of the global-batch loss (reinforce.py:105)."""
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    inst, pol, enc_in, actions, reward = _problem()
    grads, extras = _grads_on_shard("tsp", inst, pol, enc_in, actions, reward, 0, 4, 8, False)
    for g2, g1 in zip(ret["grads"], grads):
        assert torch.allclose(g2, g1, rtol=1e-4, atol=1e-7)
    assert torch.allclose(ret["extras"], extras, rtol=1e-5, atol=1e-7)
