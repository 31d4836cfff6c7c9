"""Where the end-to-end step of bench.py (public API, host buffers) spends its time.  usage: python profiles/e2e_breakdown.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import eam_rl4co_b200 as P
import bench

def timed(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): out = fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, out

def main():
    B = 1024
    dev = torch.device("cuda", 0)
    torch.backends.cuda.matmul.allow_tf32 = False
    pol = bench.make_policy().to(dev)
    env = P.TSPEnv(generator_params=dict(num_loc=bench.N_NODES), device=dev)
    host = bench.make_instances(B).pin_memory()
    t, locs = timed(lambda: host.to(dev, non_blocking=True)); print(f"H2D locs            {t:8.3f} ms")
    t, td = timed(lambda: env.reset(P.TensorDict({"locs": locs}, batch_size=[B]))); print(f"env.reset           {t:8.3f} ms")
    with torch.no_grad():
        t, (h, _) = timed(lambda: pol.encoder(td)); print(f"encoder (torch)     {t:8.3f} ms")
        t, cache = timed(lambda: pol.decoder._precompute_cache(h, 100)); print(f"cache precompute+pack {t:6.3f} ms")
        t, cache0 = timed(lambda: pol.decoder._precompute_cache(h, 0)); print(f"  of which GEMMs    {t:8.3f} ms")
    cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False)
    t, out = timed(lambda: pol._forward_fused(td, env, h, "multistart_greedy", None, 10**6, False, dict(cfg)))
    print(f"_forward_fused (precompute + alloc + rollout) {t:8.3f} ms")
    with torch.no_grad():
        t, o = timed(lambda: pol(td, env, decode_type="multistart_greedy", check=False)); print(f"policy() check=False {t:7.3f} ms")
    rows = B * 100
    hostbuf = {"reward": torch.empty(rows).pin_memory(), "log_likelihood": torch.empty(rows).pin_memory(),
               "actions": torch.empty(rows, 100, dtype=torch.int64).pin_memory()}
    def d2h():
        for k, v in hostbuf.items(): v.copy_(o[k], non_blocking=True)
    t, _ = timed(d2h); print(f"D2H outputs         {t:8.3f} ms")
    t, _ = timed(lambda: torch.zeros(rows, 100, dtype=torch.int64, device=dev)); print(f"zeros actions       {t:8.3f} ms")

if __name__ == "__main__":
    main()
