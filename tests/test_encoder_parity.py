"""CPU: the encoder mirror (eam_rl4co_b200/encoder.py, PyTorch modules) against the REFERENCE encoder's output.

tests/golden/enc_*.npz hold `h` of the unmodified reference (am/encoder.py:70-91, nn/graph/attnnet.py:16-103,
nn/ops.py:32-56, nn/env_embeddings/init.py:55-68,115-138) for synthetic weights that both sides regenerate from a
seed (oracle/synth_weights.py) and load with strict=True.  AM fixtures run eval-mode BatchNorm with non-trivial
running statistics, POMO fixtures InstanceNorm.  The `-m gpu` twin (tests/test_encoder_gpu.py) checks the tcgen05
kernels against the same `h`.  When /root/reference is present the live reference is compared as well."""
import os

import pytest
import torch

import eam_rl4co_b200 as P
from helpers import encoder_golden_files, encoder_input, load_encoder_golden

H_ATOL = 2e-5     # fp32 encoders with different summation orders (instance norm formula, BLAS): |h| is O(1..10)


@pytest.mark.parametrize("path", encoder_golden_files(), ids=lambda p: p.split("/")[-1])
def test_torch_encoder_matches_reference_fixture(path):
    torch.set_num_threads(4)
    g = load_encoder_golden(path, P)
    with torch.no_grad():
        h, _ = g["policy"].encoder(encoder_input(g["inst"]))
    err = (h - g["h"]).abs().max().item()
    print(f"{g['name']}: max |h - h_ref| = {err:.2e}, |h|max = {g['h'].abs().max().item():.2f}")
    assert h.shape == g["h"].shape and err <= H_ATOL, err


@pytest.mark.skipif(not os.path.isdir("/root/reference/rl4co"), reason="needs the reference checkout (build container)")
@pytest.mark.parametrize("env_name,pomo", [("tsp", False), ("tsp", True), ("cvrp", False), ("cvrp", True)])
def test_reference_state_dict_loads_strict_and_h_matches_live(env_name, pomo):
    from oracle import ref_runner as RR
    from oracle import rollout_oracle as O
    torch.set_num_threads(4)
    ref = RR.make_policy(env_name, pomo, seed=4321)
    mine = (P.POMOPolicy(env_name) if pomo else P.AttentionModelPolicy(env_name=env_name)).eval()
    assert list(ref.state_dict().keys()) == list(mine.state_dict().keys())
    mine.load_state_dict(ref.state_dict(), strict=True)
    torch.manual_seed(1234)
    inst = O.generate_tsp(6, 20) if env_name == "tsp" else O.generate_cvrp(6, 20)
    h_ref = RR.run_reference(env_name, inst, ref, "greedy")["hidden"].node_embeddings
    with torch.no_grad():
        h, _ = mine.encoder(encoder_input(inst))
    assert (h - h_ref).abs().max().item() <= H_ATOL
