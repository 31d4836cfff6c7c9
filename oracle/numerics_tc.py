"""TEST INFRASTRUCTURE ONLY — CPU emulation of the split-fp16 tensor-core arithmetic.

Emulates what csrc/rollout_tc.cu computes per decode step (fp16 hi/lo operand split, three
products per contraction accumulated in fp32, exp2-based masked softmax, project_out folded into
the logit key) and measures, against the oracle's own fp32 logits on a teacher-forced greedy
TSP100 POMO rollout:  max |dlogit|, and how many row-steps would pick a different greedy action.
Run:  python -m oracle.numerics_tc [B] [N]
"""
import math
import sys

import torch

from . import rollout_oracle as O


def split16(x):
    hi = x.to(torch.float16)
    lo = (x - hi.float()).to(torch.float16)
    return hi.float(), lo.float()


def mm3(a, b, acc=torch.float64):
    """a [.., M, K] x b [.., N, K]^T with both operands split into fp16 hi/lo, three products."""
    ah, al = split16(a)
    bh, bl = split16(b)
    f = lambda x, y: torch.matmul(x.to(acc), y.to(acc).transpose(-1, -2))
    return (f(ah, bh) + f(ah, bl) + f(al, bh)).float()


def tc_logits(cache, first, cur, mask, placeholder_q=None):
    """cache: folded tables [B,n,128]; first/cur [B,S]; mask [B,S,n] bool -> logits [B,S,n]."""
    B, n, E = cache["gk"].shape
    S = cur.shape[1]
    idx = lambda t, i: t.gather(1, i[..., None].expand(B, S, E))
    q = placeholder_q if placeholder_q is not None else idx(cache["pa"], first) + idx(cache["pb"], cur)
    H, d = 8, 16
    qh = q.view(B, S, H, d).transpose(1, 2)                       # [B,H,S,d]
    kh = (0.25 * cache["gk"]).view(B, n, H, d).transpose(1, 2)    # [B,H,n,d]
    vh = cache["gv"].view(B, n, H, d).transpose(1, 2)
    s = mm3(qh, kh)                                                # [B,H,S,n]
    s = s.masked_fill(~mask[:, None], float("-inf"))
    m = s.max(-1, keepdim=True).values
    p = torch.exp2((s - m) * 1.4426950408889634)
    p = torch.where(mask[:, None], p, torch.zeros_like(p))
    l = p.sum(-1, keepdim=True)
    o = mm3(p, vh.transpose(-1, -2)) / l                          # [B,H,S,d]
    heads = o.transpose(1, 2).reshape(B, S, E)
    lk = cache["lk"] / math.sqrt(E)
    return mm3(heads, lk)


def main(B=4, N=100):
    import eam_rl4co_b200 as P
    torch.manual_seed(1234)
    inst = O.generate_tsp(B, N)
    torch.manual_seed(4321)
    pol = P.POMOPolicy("tsp").eval()
    w = {k: v.detach().clone() for k, v in pol.decoder.state_dict().items()}
    with torch.inference_mode():
        h, _ = pol.encoder({"locs": inst["locs"]})
        trace = {}
        ref = O.rollout("tsp", inst, h, w, "multistart_greedy", use_graph_context=False, trace=trace)
        c = pol.decoder._precompute_cache(h)
        cache = {"gk": c.glimpse_key, "gv": c.glimpse_val, "lk": c.logit_key_folded, "pa": c.pa, "pb": c.pb}
        S = N
        acts = O.unbatchify(ref["actions"], S)                     # [B,S,T]
        first = acts[:, :, 0]
        worst, flips, total, gaps = 0.0, 0, 0, []
        for t in range(1, N):
            cur = acts[:, :, t - 1]
            mask = O.unbatchify(trace["mask"][t - 1], S)
            lg_ref = O.unbatchify(trace["logits"][t - 1], S)
            lg = tc_logits(cache, first, cur, mask)
            err = (lg - lg_ref)[mask].abs().max().item()
            worst = max(worst, err)
            x = torch.tanh(lg) * 10
            x = x.masked_fill(~mask, float("-inf"))
            a = x.argmax(-1)
            diff = a != acts[:, :, t]
            flips += int(diff.sum())
            total += diff.numel()
            if diff.any():
                xr = (torch.tanh(lg_ref) * 10).masked_fill(~mask, float("-inf"))
                top2 = xr.topk(2, -1).values
                gaps += (top2[..., 0] - top2[..., 1])[diff].tolist()
        print(f"TSP{N} POMO B={B}: max|dlogit|={worst:.3e}  flipped row-steps={flips}/{total}  "
              f"oracle top-2 gaps at flips={['%.2e' % g for g in gaps[:10]]}")


if __name__ == "__main__":
    main(*(int(a) for a in sys.argv[1:]))
