#!/usr/bin/env python
"""bench.py — rollout instance-steps/s on the BASELINE.json headline workload.

Workload (config c4 of BASELINE.json): TSP100, POMO policy (6 encoder layers, instance norm, no graph
context), multistart greedy with S=100 starts per instance, B=1024 instances PER GPU (weak scaling:
instances are independent, each rank rolls out its own shard, no data-path collective).
One "step" = one full rollout of the batch = B*S rows x T=100 actions = 1.024e7 instance-steps.

  value : whole-job instance-steps/s, inputs (encoder output h, locs, weights) resident in HBM;
          timed region = cache precompute GEMMs + operand pack kernel + the persistent tcgen05 rollout
          kernel, per step.
  e2e   : the same metric through the public API `policy(td, env, decode_type=...)` with HOST buffers:
          pinned host locs -> H2D -> env.reset -> encoder -> rollout -> D2H of reward / log_likelihood /
          actions, every step.
  roofline : the rollout kernel alone (CUDA events around its launch on the launching stream) against
          SURVEY.md §8(d)'s algorithmic bytes per instance-step and MEASURED_PEAKS.json's HBM figure.
  --impl reference : the reference's CPU implementation of the path (oracle port of the unmodified
          reference, bit-identical to it on CPU — see oracle/), all host threads, bounded sample.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

N_NODES = 100
NUM_STARTS = 100
B_PER_GPU = 1024
ALGO_BYTES_PER_INSTANCE_STEP = 2302
METRIC = "rollout_instance_steps_per_sec"
UNIT = "instance-steps/s"
WORKLOAD = "TSP100 POMO multistart greedy, S=100, B=1024 per GPU, random-init weights"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU, help="instances per GPU")
    ap.add_argument("--cpu-sample", type=int, default=128, help="instances in the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_traffic():
    """DRAM bytes per launch of the rollout kernel from the committed `ncu --set full` capture."""
    p = os.path.join(ROOT, "profiles", "r01_rollout_tc_traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get("dram_bytes_per_launch")
    return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i] == "Active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_instances(B, seed=1234):
    from eam_rl4co_b200 import TSPGenerator
    torch.manual_seed(seed)
    return TSPGenerator(num_loc=N_NODES)(B)["locs"]            # host, [B, 100, 2]


def make_policy():
    import eam_rl4co_b200 as P
    torch.manual_seed(4321)
    return P.POMOPolicy("tsp").eval()


# ---------------------------------------------------------------------------------------------
# CPU leg: the reference's implementation of the path (oracle port) on the host cores
# ---------------------------------------------------------------------------------------------
def cpu_reference_run(sample_B, steps, warmup):
    """Times oracle.rollout (decode loop given h: cache precompute + T decoder/selection/env steps +
    reward), all host threads, on `sample_B` instances of the same workload.  Returns dict."""
    from oracle import rollout_oracle as O
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    pol = make_policy()
    w = {k: v.detach().clone() for k, v in pol.decoder.state_dict().items()}
    locs = make_instances(sample_B)
    with torch.inference_mode():
        h, _ = pol.encoder({"locs": locs})
        times = []
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            out = O.rollout("tsp", {"locs": locs}, h, w, "multistart_greedy", use_graph_context=False)
            dt = time.perf_counter() - t0
            if it >= warmup:
                times.append(dt)
    isteps = out["actions"].numel()
    total = sum(times)
    return {"value": isteps * len(times) / total, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{sample_B} instances x {NUM_STARTS} starts x {N_NODES} steps per rollout, "
                      f"{len(times)} rollouts, oracle port of the reference (torch CPU, {cores} threads)",
            "ms_per_step": 1e3 * total / len(times), "instance_steps_per_step": isteps}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample: 128 instances per rollout is the CPU path's best operating point on these hosts (3.2e6
    # instance-steps/s; 1.4e6 at 16 instances, 2.0e6 at 512), ~0.4 s per rollout
    steps = max(1, min(args.steps, 50))
    warm = max(1, min(args.warmup, 2))
    r = cpu_reference_run(args.cpu_sample, steps, warm)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# B200 leg
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist
    import eam_rl4co_b200 as P
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the rollout path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # keep stdout to the one JSON line: NCCL's version banner goes there at NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    B, K, W = args.batch, args.steps, max(args.warmup, 3)

    pol = make_policy().to(dev)
    env = P.TSPEnv(generator_params=dict(num_loc=N_NODES), device=dev)
    # each rank owns its own shard of instances (global instance index = rank*B + b keys the noise)
    locs_host = make_instances(B, seed=1234 + rank).pin_memory()
    locs_dev = locs_host.to(dev)
    td_dev = env.reset(P.TensorDict({"locs": locs_dev}, batch_size=[B]))
    with torch.no_grad():
        h, _ = pol.encoder(td_dev)
    rows = B * NUM_STARTS
    isteps = rows * N_NODES
    events = []
    cfg = dict(temperature=1.0, tanh_clipping=10.0, mask_logits=True, store_all_logp=False,
               instance_offset=rank * B, _events=events)

    def device_step():
        # hot path with inputs resident in HBM: cache precompute + ONE persistent rollout kernel
        return pol._forward_fused(td_dev, env, h, "multistart_greedy", None, 10**6, False, cfg)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(W):
        device_step()
    barrier()
    events.clear()
    # one sampler per job (rank 0's GPU): eight nvidia-smi pollers on one host contend with the CUDA driver for the
    # same locks and show up in the end-to-end leg
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t0.record()
    for _ in range(K):
        device_step()
    t1.record()
    barrier()
    ms_total = t0.elapsed_time(t1)
    kernel_ms = [a.elapsed_time(b) for a, b in events]

    # ---- e2e: public API, host buffers, H2D + D2H inside the timed region --------------------
    out_host = {"reward": torch.empty(rows, dtype=torch.float32).pin_memory(),
                "log_likelihood": torch.empty(rows, dtype=torch.float32).pin_memory(),
                "actions": torch.empty(rows, N_NODES, dtype=torch.int64).pin_memory()}
    h2d = locs_host.numel() * 4
    d2h = sum(v.numel() * v.element_size() for v in out_host.values())

    # result copies run on their own stream so that the D2H of step i overlaps the compute of step i+1; the timed
    # region ends with a full device synchronize, so every copy is inside it
    copy_stream = torch.cuda.Stream(device=dev)

    def e2e_step():
        td = P.TensorDict({"locs": locs_host.to(dev, non_blocking=True)}, batch_size=[B])
        td = env.reset(td)
        with torch.no_grad():
            out = pol(td, env, decode_type="multistart_greedy", check=False, instance_offset=rank * B)
        copy_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(copy_stream):
            for k, v in out_host.items():
                v.copy_(out[k], non_blocking=True)
                out[k].record_stream(copy_stream)
        return out

    for _ in range(W):
        e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    torch.cuda.current_stream().wait_stream(copy_stream)     # the last step's copies are part of the timed region
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None   # sampled over both timed regions (device-resident and end-to-end)

    from eam_rl4co_b200 import sharding
    k_mean = statistics.mean(kernel_ms)
    ms_total, e2e_ms, k_mean = sharding.max_over_ranks([ms_total, e2e_ms, k_mean], device=dev)   # device-timed, MAX
    if rank == 0:
        peak, peak_src = peaks()
        k_ms = k_mean
        achieved = ALGO_BYTES_PER_INSTANCE_STEP * isteps / (k_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": world * isteps * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "instances_per_gpu": B, "num_starts": NUM_STARTS,
                       "nodes": N_NODES, "decode_type": "multistart_greedy", "parallelism": f"dp{world}",
                       "instance_steps_per_step_per_gpu": isteps,
                       "l2": "inputs larger than L2 (tables 262 MB + outputs 123 MB per step)",
                       "timed": "cache precompute GEMMs + operand pack + persistent tcgen05 rollout kernel"},
            "e2e": {"value": world * isteps * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / K,
                    "includes": "H2D locs, env.reset, encoder (tcgen05 kernels: 4 GEMMs + attention per layer), "
                                "cache precompute, rollout kernel, D2H reward+log_likelihood+actions"},
            "gpu_launches": 2 * len(kernel_ms),      # per step: r4::tc::tc_pack_kernel + r4::tc::rollout_tc_kernel
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": measured_traffic(), "peak_source": peak_src,
                         "kernel": "r4::tc::rollout_tc_kernel<TSP, GREEDY>", "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": ALGO_BYTES_PER_INSTANCE_STEP * isteps,
                         "note": "algorithmic bytes per SURVEY.md §8(d) (K,V,L re-read every step); the "
                                 "kernel keeps the operand tiles in shared memory for the whole rollout, so "
                                 "DRAM traffic (`traffic`, ncu) is ~60x lower and the binding resources are "
                                 "on-chip: CUDA-core issue + MUFU for the softmax/selection around the "
                                 "tcgen05 MMAs (DESIGN.md, profiles/)"},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference_run(args.cpu_sample, 30, 1)      # ~12 s of CPU work
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
