#!/usr/bin/env python
"""rover env-steps/s of the batched stepping path (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload config2|config3|config4|config5]

A "step" is one pass of the hot path over one batch: tfode_batch_step(SUBSTEPS) for every environment of the
workload (SUBSTEPS x { Environment::next_step; Rover_1::next_step } per env, one fused kernel launch).
N=1 workload (default) = BASELINE.json configs[1]: rover_1_mt, 4096 rovers on the rock-groove heightfield,
random actions every 0.5 s.  N>1: every rank owns 4096 more environments (weak scaling, no data-path collective).
`--impl reference` times the CPU path instead: the oracle port (the reference itself cannot be built here: libode,
Eigen, Boost are absent -- DESIGN.md), all host threads, on a bounded sample of the same workload.
Every arm -- device-resident `value`, host-buffer `e2e`, the CPU arms -- starts from a reset and times the same episode
segment (sub-steps W*100 .. (W+K)*100): a sub-step's cost drifts by +-20 % along the episode (19.3 ms per 100 sub-steps
on the plane before the terrain, 28 ms with all wheels on the groove; tools/e2e_split.py), so arms timed on different
segments are not comparable.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALGO_BYTES_PER_ENV_STEP = 1680  # (read + write) x 840 B of persistent fp32 state
SUBSTEPS = 100                  # env-steps per env per bench step (one launch)
CPU_SAMPLE_ENVS = 4096          # the CPU arm steps at most this many environments of the workload

WORKLOADS = {
    # name: (description, envs per GPU)
    "config2": ("rover_1_mt: 4096 rovers, rock-groove heightfield 200x200 at (1,0,-0.3), random actions every 0.5 s", 4096),
    "config3": ("rover_training_1: 16384 rovers, step boxes, fused 17-512-512-2 policy every 0.5 s, e-greedy", 16384),
    "config4": ("rover_1_crawlers: 8192 rovers per GPU, plane, Crawler_1 gait with per-env jitter", 8192),
    "config5": ("rock_groove_large sweep: 131072 rovers per GPU, 400x400 heightfield 6x6 m, random actions", 131072),
}


def make_cfg(workload, n_envs, env_id_offset=0, device=0):
    from tf_ode_b200 import abi, make_config, reference_heightmap
    kw = dict(env_id_offset=env_id_offset, device=device)
    if workload == "config2":
        hm = reference_heightmap("rock_groove")  # the reference's env_data/heightmap_rock_groove.png (levels fixture)
        return make_config(abi.PRESET_MT_RANDOM, n_envs, heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3)), **kw)
    if workload == "config3":
        return make_config(abi.PRESET_TRAINING, n_envs, policy_precision=abi.POLICY_TC, **kw)  # actor on tcgen05 (the CPU arm runs it in fp64)
    if workload == "config4":
        return make_config(abi.PRESET_CRAWLING, n_envs, crawler_amp=25.0, crawler_amp_jitter=5.0, crawler_rate_jitter=5.0, **kw)
    if workload == "config5":
        hm = reference_heightmap("rock_groove_large")
        return make_config(abi.PRESET_MT_RANDOM, n_envs, heightfield=(hm, 6.0, 6.0, (2.0, 0.0, -0.3)), **kw)
    raise SystemExit("unknown workload %r" % workload)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def mark(self):
        """Start of the timed region: samples written before this point (the warm-up steps, same load) are used only if nvidia-smi
        delivers none after it (a 0.2 s timed region on an 8-GPU box can end before its first 100 ms tick is written)."""
        try:
            self.skip = sum(1 for _ in open(self.f.name))
        except OSError:
            self.skip = 0

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        rows = [r.split(",") for r in self.f.read().strip().splitlines() if r.count(",") >= 6]
        os.unlink(self.f.name)
        skip = getattr(self, "skip", 0)
        if len(rows) > skip:
            rows = rows[skip:]
        elif rows:
            out["note"] = "sampled during the warm-up steps (same load); the timed region ended before nvidia-smi's next tick"
        if not rows:
            return out
        sm = sorted(float(r[0]) for r in rows)
        pw = [float(r[2]) for r in rows]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names) if any("Active" in r[3 + i] and "Not" not in r[3 + i] for r in rows)]
        out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=float(rows[0][1]), reasons=reasons, power_w_max=max(pw), samples=len(rows))
        return out


def action_tape(rank, k, n_envs):
    """Step k's externally supplied actions of rank `rank`: U(-15,15) deg/s, U(-25,25) N.m per environment.  Every arm feeds the same tape
    (they act until a scene's own controller takes over at its first IC tick), so every arm simulates the same trajectories."""
    import numpy as np
    return (np.random.RandomState(100003 * rank + k).uniform(-1, 1, size=(n_envs, 2)) * np.array([15.0, 25.0])).astype(np.float32)


def cpu_segment(workload, n_envs, warmup, steps, max_env_steps=None):
    """oracle<double> over all host threads on the GPU arm's own episode segment: from reset, `warmup` untimed then `steps` timed steps of
    SUBSTEPS sub-steps (the cost of a sub-step drifts by +-20 % along the episode as the rovers reach the terrain, so the arms must cover
    the same sub-steps).  Bounded sample: the first min(n_envs, CPU_SAMPLE_ENVS) environments, fewer if max_env_steps says so."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from tf_ode_b200 import abi, glorot_policy
    cores = os.cpu_count() or 1
    n = min(n_envs, CPU_SAMPLE_ENVS)
    if max_env_steps:
        n = max(256, min(n, int(max_env_steps // ((warmup + steps) * SUBSTEPS)) // 256 * 256))
    cfg = make_cfg(workload, n)
    o = oracle_lib.Oracle(cfg, "double", threads=cores)
    if cfg.controller == abi.CTRL_MLP:
        o.load_policy(glorot_policy())
    tape = [action_tape(0, k, n_envs)[:n] for k in range(warmup + steps)]
    for k in range(warmup):
        o.set_actions(tape[k]); o.step(SUBSTEPS)
    t0 = time.perf_counter()
    for k in range(warmup, warmup + steps):
        o.set_actions(tape[k]); o.step(SUBSTEPS)
    dt = time.perf_counter() - t0
    o.close()
    sample = ("%d envs x %d sub-steps per step, episode sub-steps %d..%d from reset (same scene/controller/segment as the GPU arm), oracle<double> "
              "(reference-equivalent fp64 restatement; real libode unavailable on this box), %d std::threads"
              % (n, SUBSTEPS, warmup * SUBSTEPS, (warmup + steps) * SUBSTEPS, cores))
    return n, dt, cores, sample


def run_reference(args, n_envs):
    """The CPU arm: oracle<double> (reference-equivalent arithmetic), N envs over all host threads."""
    n_envs, dt, cores, sample = cpu_segment(args.workload, n_envs, args.warmup, args.steps)
    sub = SUBSTEPS
    v = n_envs * sub * args.steps / dt
    return {
        "impl": "reference", "metric": "rover env-steps/sec", "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": args.workload + ": " + WORKLOADS[args.workload][0], "envs": n_envs, "substeps_per_step": sub},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="environments per GPU (default: the workload's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    n_envs = args.envs or WORKLOADS[args.workload][1]

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args, n_envs)))
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from tf_ode_b200 import RoverBatch, glorot_policy, abi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the stepping path has no CPU implementation")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"   # the version banner goes to stdout, which carries the one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfg = make_cfg(args.workload, n_envs, env_id_offset=rank * n_envs, device=local)
    g = RoverBatch(cfg, stream=torch.cuda.current_stream().cuda_stream)
    if cfg.controller == abi.CTRL_MLP:
        g.load_policy(glorot_policy())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    # -------- device-resident arm: state already in HBM, one fused launch per step ------------------------------------------------
    tape = [action_tape(rank, k, n_envs) for k in range(args.warmup + args.steps)]   # the same inputs for every arm
    tape_dev = [torch.from_numpy(t).cuda() for t in tape]
    g.reset()
    clocks = ClockSampler(local)  # started before the warm-up so that nvidia-smi is up when the timed region begins
    for k in range(args.warmup):
        g.set_actions(tape_dev[k]); g.step(SUBSTEPS)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    clocks.mark()
    l0 = g.launch_count
    t_wall = time.perf_counter()
    for k, (a, b) in enumerate(ev):
        flush.fill_(1)  # L2 flush between timed iterations, outside the event pair
        a.record()
        g.set_actions(tape_dev[args.warmup + k])  # device-resident inputs: set_actions_kernel, counted in gpu_launches
        g.step(SUBSTEPS)
        b.record()
    barrier()
    t_wall = time.perf_counter() - t_wall
    launches = g.launch_count - l0
    clk = clocks.stop()
    ms = [a.elapsed_time(b) for a, b in ev]
    tot_ms = torch.tensor([sum(ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot_ms, op=dist.ReduceOp.MAX)
    tot_ms = float(tot_ms.item())
    units = n_envs * SUBSTEPS * args.steps * world
    value = units / (tot_ms * 1e-3)
    step_launches = launches / args.steps  # set_actions_kernel + step kernel (+ policy kernels for the MLP workload)

    # -------- end-to-end arm: HOST buffers through the C-ABI round trip (H2D actions, step, D2H obs/reward/done) --------------------
    act = torch.zeros(n_envs, 2, dtype=torch.float32).pin_memory().numpy()
    obs = torch.zeros(n_envs, 17, dtype=torch.float32).pin_memory().numpy()
    rew = torch.zeros(n_envs, dtype=torch.float32).pin_memory().numpy()
    done = torch.zeros(n_envs, dtype=torch.uint8).pin_memory().numpy()
    g.reset()  # same episode segment and inputs as the device-resident arm: sub-steps W*100 .. (W+K)*100 from reset
    for k in range(args.warmup):
        act[:] = tape[k]
        g.rollout_host(act, SUBSTEPS, obs, rew, done)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        act[:] = tape[args.warmup + k]  # this step's inputs reach the pinned buffer inside the timed region
        g.rollout_host(act, SUBSTEPS, obs, rew, done)  # actions are uploaded every step; scenes with their own controller overwrite them at IC ticks
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = units / float(e2e_t.item())

    # -------- CPU baseline beside it (rank 0, N=1 only): bounded sample of the same workload ------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cn, dtc, cores, sample = cpu_segment(args.workload, n_envs, args.warmup, args.steps, max_env_steps=12e6)
        cpu = {"value": cn * SUBSTEPS * args.steps / dtc, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        kern_ms = tot_ms / args.steps
        traffic = None  # DRAM bytes per launch of the committed ncu capture (same workload, same envs/sub-steps per launch)
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload)
            if t and t["envs"] == n_envs and t["substeps"] == SUBSTEPS:
                traffic = t["dram_bytes_per_launch"]
        except Exception:
            pass
        achieved = ALGO_BYTES_PER_ENV_STEP * n_envs * SUBSTEPS / (kern_ms * 1e-3) / 1e9
        line = {
            "metric": "rover env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload + ": " + WORKLOADS[args.workload][0], "envs_per_gpu": n_envs, "substeps_per_step": SUBSTEPS,
                       "pgs_iters": cfg.pgs_iters, "dt": 0.001, "sharding": "contiguous env ranges, no data-path collective",
                       "l2": "256 MiB fill between timed steps, outside the per-step CUDA-event pairs",
                       "segment": "every arm (value, e2e, CPU) starts from reset and times episode sub-steps %d..%d" % (args.warmup * SUBSTEPS, (args.warmup + args.steps) * SUBSTEPS)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read+write)", "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                         "kernel": "tfode::step_kernel", "kernel_ms": kern_ms,
                         "note": "algorithmic bytes = 1680 B per env-step (SURVEY 8(d)); the path is FP32-issue/latency bound (sequential PGS), see DESIGN.md"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(act.nbytes) * world,
                    "d2h_bytes_per_step": int(obs.nbytes + rew.nbytes + done.nbytes) * world},
            "gpu_launches": int(launches), "launches_per_step": step_launches, "clocks": clk, "wall_s_timed_region": t_wall,
        }
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    g.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
