#!/usr/bin/env python
"""rover env-steps/s of the batched stepping path (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload config2|config3|config4|config5]
                  [--envs E] [--no-verify] [--no-cpu-baseline]

A "step" is one pass of the hot path over one batch: tfode_batch_step(S) for every environment of the workload, i.e.
S x { Environment::next_step; Rover_1::next_step } per environment.  S = 500 sub-steps (one command period of the rover,
SetCmdPeriod(0.5)) for configs 2-4 and 100 for config 5, so the default K = 20 timed steps cover the 10 000 (2 000)
sub-steps SURVEY 8(d) asks for.  Every arm (device-resident `value`, host-buffer `e2e`, the CPU arms, the parity check)
runs W untimed warm-up steps, resets, and times/steps episode sub-steps 0 .. K*S from that reset, so all arms cover the
same segment of the episode (the cost of a sub-step drifts by +-20 % along it).

Workloads (BASELINE.json configs):
  N = 1 default  config3  rover_training_1: 16384 rovers, step boxes, 17-512-512-2 actor fused into the step kernel
  N > 1 default  config5  rock_groove_large sweep: 2^20 rovers in total, 2^20 / N per GPU (strong scaling)
  --workload config2 (4096 rovers per GPU, rock-groove heightfield) and config4 (8192 crawlers per GPU) select the others.
`--impl reference` times the CPU path instead: the oracle port (the reference itself cannot be built here: libode, Eigen,
Boost are absent -- DESIGN.md), all host threads, on a bounded sample of the same workload and segment.
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
CPU_SAMPLE_ENV_STEPS = 12e6     # the CPU legs step about this many env-steps (about 20 s on 16 host threads)
CPU_SAMPLE_MAX_ENVS = 4096      # ... and at most this many environments
VERIFY_ENVS = 16                # environments re-run on oracle<float> after the timed arms (bench.py --verify, on by default)
TOTAL_ENVS_CONFIG5 = 1 << 20

WORKLOADS = {
    # name: (description, envs per GPU, sub-steps per bench step)
    "config2": ("rover_1_mt: 4096 rovers, rock-groove heightfield 200x200 at (1,0,-0.3), random actions every 0.5 s", 4096, 500),
    "config3": ("rover_training_1: 16384 rovers, step boxes, 17-512-512-2 actor every 0.5 s fused into the step kernel, e-greedy", 16384, 500),
    "config4": ("rover_1_crawlers: 8192 rovers per GPU, plane, Crawler_1 gait with per-env jitter", 8192, 500),
    "config5": ("rock_groove_large sweep: 2^20 rovers in total (2^20 / N per GPU), 400x400 heightfield 6x6 m, random actions", TOTAL_ENVS_CONFIG5, 100),
}
SUBSTEPS = 100  # developer tools (tools/ab_bench.py, tools/phase_timing.py) step in units of 100


def dev_envs(workload):
    """environments per GPU the developer tools use (config 5: 131072, an eighth of the sweep)"""
    return TOTAL_ENVS_CONFIG5 // 8 if workload == "config5" else WORKLOADS[workload][1]


SOLVER = 0   # set from --solver: every arm (GPU, CPU, verify) builds its configuration through make_cfg


def make_cfg(workload, n_envs, env_id_offset=0, device=0, **over):
    from tf_ode_b200 import abi, make_config, reference_heightmap
    kw = dict(env_id_offset=env_id_offset, device=device, solver=SOLVER)
    kw.update(over)
    if workload == "config2":
        hm = reference_heightmap("rock_groove")  # the reference's env_data/heightmap_rock_groove.png
        return make_config(abi.PRESET_MT_RANDOM, n_envs, heightfield=(hm, 3.0, 3.0, (1.0, 0.0, -0.3)), **kw)
    if workload == "config3":
        return make_config(abi.PRESET_TRAINING, n_envs, **kw)  # policy_precision FP32: the actor runs fused inside step_kernel
    if workload == "config4":
        return make_config(abi.PRESET_CRAWLING, n_envs, crawler_amp=25.0, crawler_amp_jitter=5.0, crawler_rate_jitter=5.0, **kw)
    if workload == "config5":
        hm = reference_heightmap("rock_groove_large")
        return make_config(abi.PRESET_MT_RANDOM, n_envs, heightfield=(hm, 6.0, 6.0, (2.0, 0.0, -0.3)), **kw)
    raise SystemExit("unknown workload %r" % workload)


def config_block(workload, n_envs, world, substeps, steps):
    """The `config` object: identical keys and values in the GPU arm's and the reference arm's line."""
    from tf_ode_b200 import make_config, abi
    return {"workload": workload + ": " + WORKLOADS[workload][0], "envs_per_gpu": n_envs, "envs_total": n_envs * world, "substeps_per_step": substeps,
            "pgs_iters": make_config(abi.PRESET_SCENE_1, 1).pgs_iters, "dt": 0.001, "solver": "dantzig (dWorldStep equivalent)" if SOLVER else "quickstep",
            "sharding": "contiguous env ranges, no collective inside the step; NCCL all_gather of the rollout rows per step in the N>1 e2e arm",
            "l2": "256 MiB fill between timed steps, outside the per-step CUDA-event pairs (GPU arm)",
            "segment": "every arm starts from a reset and covers episode sub-steps 0..%d" % (steps * substeps)}


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
        delivers none after it."""
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


def action_tape(rank, k, n_envs, first=0):
    """Step k's externally supplied actions for global environments [first, first + n_envs): U(-15,15) deg/s, U(-25,25) N.m, a
    function of (k, global id) only, so every arm -- and every sharding of the same job -- feeds each rover the same commands
    (they act until a scene's own controller takes over at its first IC tick)."""
    import numpy as np
    g = np.arange(first, first + n_envs, dtype=np.uint64)
    out = np.empty((n_envs, 2), np.float32)
    for j, scale in enumerate((15.0, 25.0)):
        x = (g * np.uint64(2654435761) + np.uint64(k * 40503 + j * 9176 + 12345)) & np.uint64(0xFFFFFFFF)
        x ^= x >> np.uint64(15); x = (x * np.uint64(2246822519)) & np.uint64(0xFFFFFFFF)
        x ^= x >> np.uint64(13); x = (x * np.uint64(3266489917)) & np.uint64(0xFFFFFFFF)
        x ^= x >> np.uint64(16)
        out[:, j] = ((x >> np.uint64(8)).astype(np.float64) / 16777216.0 * 2.0 - 1.0) * scale
    return out


def oracle_mod():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    return oracle_lib


def cpu_segment(workload, n_envs, substeps, steps, precision="double", threads=None, budget=CPU_SAMPLE_ENV_STEPS):
    """The CPU path on the GPU arm's own episode segment: from reset, `steps` timed steps of `substeps` sub-steps on the first
    n environments of the workload, n bounded so that the leg stays around `budget` env-steps.  Returns (n, seconds, threads, sample text)."""
    oracle_lib = oracle_mod()
    from tf_ode_b200 import abi, glorot_policy
    cores = threads or os.cpu_count() or 1
    n = int(budget // (steps * substeps)) // 64 * 64
    n = max(64, min(n, n_envs, CPU_SAMPLE_MAX_ENVS))
    cfg = make_cfg(workload, n)
    o = oracle_lib.Oracle(cfg, precision, threads=cores)
    if cfg.controller == abi.CTRL_MLP:
        o.load_policy(glorot_policy())
    o.set_actions(action_tape(0, 0, n)); o.step(10)   # page-in
    o.reset()
    t0 = time.perf_counter()
    for k in range(steps):
        o.set_actions(action_tape(0, k, n)); o.step(substeps)
    dt = time.perf_counter() - t0
    o.close()
    kind = "oracle<%s>" % precision
    sample = ("first %d envs of the workload x %d sub-steps per step x %d steps, episode sub-steps 0..%d from reset (same scene / controller / action "
              "tape / segment as the GPU arm), %s (%s restatement; real libode unavailable on this box), %d std::threads"
              % (n, substeps, steps, steps * substeps, kind, "reference-equivalent fp64" if precision == "double" else "GPU-equivalent fp32", cores))
    return n, dt, cores, sample


def run_reference(args, n_envs, world, substeps):
    """The CPU arm: oracle<double> (reference-equivalent arithmetic), a bounded sample of the workload over all host threads."""
    n, dt, cores, sample = cpu_segment(args.workload, n_envs, substeps, args.steps)
    v = n * substeps * args.steps / dt
    return {
        "impl": "reference", "metric": "rover env-steps/sec", "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": scaling_of(args.workload), "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": config_block(args.workload, n_envs, world, substeps, args.steps),
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "per_core": v / cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def scaling_of(workload):
    return "strong" if workload == "config5" else "weak"


def verify_against_oracle(g, workload, n_envs, first, substeps, warmup, steps, ids):
    """bench.py --verify: the sampled environments (local ids) of the batch `g`, which has just run `warmup` steps, a reset and `steps`
    steps over the action tape, re-run one by one on oracle<float> through the same calls; bodies, LCG state, contact count, episode
    bookkeeping must be bit-equal.  Returns (checked, bit_exact, first mismatch text)."""
    import numpy as np
    oracle_lib = oracle_mod()
    from tf_ode_b200 import abi, glorot_policy
    bad = None
    for e in ids:
        cfg = make_cfg(workload, 1, env_id_offset=first + e)
        o = oracle_lib.Oracle(cfg, "float", threads=1)
        if cfg.controller == abi.CTRL_MLP:
            o.load_policy(glorot_policy())
        for k in range(warmup):
            o.set_actions(action_tape(0, k, 1, first + e)); o.step(substeps)
        o.reset()
        for k in range(steps):
            o.set_actions(action_tape(0, k, 1, first + e)); o.step(substeps)
        sg, so = g.get_state(e), o.get_state(0)
        nb = 12 if cfg.scene == abi.SCENE_STEP_BOXES else 10
        bg, bo = oracle_lib.state_bodies(sg)[:nb], oracle_lib.state_bodies(so)[:nb]
        ints = ("lcg_seed", "n_contacts", "tick", "time_steps", "episode", "episodes_done")
        ok = np.array_equal(bg, bo) and all(getattr(sg, f) == getattr(so, f) for f in ints) and sg.total_reward == so.total_reward
        o.close()
        if not ok and bad is None:
            bad = "env %d: max |d body| = %.3e, %s" % (first + e, float(np.abs(bg - bo).max()), {f: (getattr(sg, f), getattr(so, f)) for f in ints})
    return len(ids), bad is None, bad


def actor_tc_line(peaks):
    """The batched tensor-core actor (policy_tc_kernel: tcgen05 / TMEM / bulk-TMA) on 16384 rows through tfode_policy_infer, timed with
    CUDA events two ways: a CUDA graph of 50 launches replayed back to back (the kernel's rate: one ctypes call costs the host as much as
    the kernel runs, so a Python loop of launches would time the host), and ONE launch bracketed by events (its latency, launch overhead
    included).  Tensor-pipe roofline of the first against the measured bf16 peak."""
    import numpy as np
    import torch
    from tf_ode_b200 import RoverBatch, abi, glorot_policy, make_config
    rows, per_graph, replays = 16384, 50, 40
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        g = RoverBatch(make_config(abi.PRESET_TRAINING, 8, policy_precision=abi.POLICY_TC), stream=side.cuda_stream)
        g.load_policy(glorot_policy())
        x = torch.from_numpy(np.random.RandomState(3).rand(rows, 17).astype(np.float32) * 2 - 1).cuda()
        y = torch.zeros(rows, 2, device="cuda")
        for _ in range(5):
            g.policy_infer(x, y)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
        for a, b in ev:
            a.record(side); g.policy_infer(x, y); b.record(side)
        side.synchronize()
        single = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            for _ in range(per_graph):
                g.policy_infer(x, y)
        graph.replay(); side.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(side)
        for _ in range(replays):
            graph.replay()
        b.record(side)
        side.synchronize()
        us = a.elapsed_time(b) * 1e3 / (replays * per_graph)
        del graph
        g.close()
    flops = 2.0 * rows * (32 * 512 + 512 * 512)   # the MMAs the kernel issues (layer 1 padded to K = 32) -- the 512 -> 2 output layer is CUDA-core work
    peak = float(peaks.get("bf16_tflops", 1668.0))
    tf = flops / (us * 1e-6) / 1e12
    return {"kernel": "tfode::policy_tc_kernel", "rows": rows, "us_per_launch": us, "launches_timed": replays * per_graph,
            "how": "CUDA graph of %d tfode_policy_infer launches replayed %d times back to back, one CUDA-event pair around all of them" % (per_graph, replays),
            "us_single_launch_median": single[len(single) // 2], "us_single_launch_min": single[0], "tflops": tf,
            "roofline": {"bound": "tensor", "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                         "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst: the kernel is timed alone)" if peaks else "fallback"},
            "note": "batched tfode_policy_infer; inside tfode_batch_step the fp32 actor runs fused in step_kernel (no parking), which is faster end to end"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="environments per GPU (default: the workload's)")
    ap.add_argument("--substeps", type=int, default=0, help="sub-steps per bench step (default: the workload's)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-actor-tc", action="store_true")
    ap.add_argument("--solver", default="quickstep", choices=["quickstep", "dantzig"], help="world step: dWorldQuickStep (default) or the dWorldStep-equivalent dense Dantzig LCP")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    global SOLVER
    SOLVER = 1 if args.solver == "dantzig" else 0
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    nominal_world = max(world, args.gpus)
    if args.workload is None:
        args.workload = "config3" if nominal_world == 1 else "config5"
    n_envs = args.envs or (TOTAL_ENVS_CONFIG5 // nominal_world if args.workload == "config5" else WORKLOADS[args.workload][1])
    substeps = args.substeps or WORKLOADS[args.workload][2]

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(run_reference(args, n_envs, nominal_world, substeps)))
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    from tf_ode_b200 import RoverBatch, glorot_policy, abi
    from tf_ode_b200.rollout import ShardedRollout

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

    first = rank * n_envs
    n_total = n_envs * world
    roll = ShardedRollout(lambda n, f, dev: make_cfg(args.workload, n, env_id_offset=f, device=dev), n_total, device=local)
    g = roll.batch
    assert (roll.first, roll.count) == (first, n_envs)
    if g.cfg.controller == abi.CTRL_MLP:
        g.load_policy(glorot_policy())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    W, K = args.warmup, args.steps

    # -------- device-resident arm: state and inputs already in HBM, one tfode_batch_step per step -------------------------------------
    tape = [action_tape(rank, k, n_envs, first) for k in range(max(W, K))]   # the same inputs for every arm
    tape_dev = [torch.from_numpy(t).cuda() for t in tape]
    clocks = ClockSampler(local)  # started before the warm-up so that nvidia-smi is up when the timed region begins
    for k in range(W):
        g.set_actions(tape_dev[k]); g.step(substeps)
    g.reset()
    g.contact_stats(reset=True)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    clocks.mark()
    l0 = g.launch_count
    t_wall = time.perf_counter()
    for k, (a, b) in enumerate(ev):
        flush.fill_(1)  # L2 flush between timed iterations, outside the event pair
        a.record()
        g.set_actions(tape_dev[k])  # device-resident inputs: set_actions_kernel, counted in gpu_launches
        g.step(substeps)
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
    units = n_envs * substeps * K * world
    value = units / (tot_ms * 1e-3)

    # -------- parity of the timed batch itself (--verify): sampled environments re-run on oracle<float>, contact-capacity audit -----------
    parity = None
    over, max_gen = g.contact_stats()
    if not args.no_verify:
        per_rank = max(2, VERIFY_ENVS // world)
        ids = sorted(set(int(x) for x in np.random.RandomState(977 + rank).randint(0, n_envs, size=per_rank)) | {0, n_envs - 1})[:per_rank + 2]
        checked, ok, why = verify_against_oracle(g, args.workload, n_envs, first, substeps, W, K, ids)
        stat = torch.tensor([checked, 0 if ok else 1, over, max_gen], dtype=torch.int64, device="cuda")
        if world > 1:
            mx = stat.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            dist.all_reduce(stat, op=dist.ReduceOp.SUM); stat[3] = mx[3]
        parity = {"envs_checked": int(stat[0]), "bit_exact": int(stat[1]) == 0,
                  "what": "bodies (pos, quat, lin/ang velocity), lcg_seed, n_contacts, tick, time_steps, episode counters, total_reward of the sampled "
                          "environments of the TIMED batch after its last step == oracle<float> run through the same warm-up, reset and action tape",
                  "contact_overflow_steps": int(stat[2]), "max_generated_contacts": int(stat[3]), "max_contacts": int(g.cfg.max_contacts)}
        if why:
            parity["first_mismatch"] = why

    # -------- end-to-end arm: HOST buffers.  N = 1: the C-ABI round trip tfode_batch_rollout_host (H2D actions, step, D2H obs / reward / done).
    # N > 1: each rank uploads its actions, steps, packs its rollout rows on the device; ONE NCCL all_gather per step (second stream, overlapped
    # with the next step) brings every rank's rows to the trainer rank, which reads them back to pinned host memory.
    act = torch.zeros(n_envs, 2, dtype=torch.float32).pin_memory()
    act_np = act.numpy()
    for k in range(W):
        g.set_actions(tape_dev[k]); g.step(substeps)
    g.reset()
    if world == 1:
        obs = torch.zeros(n_envs, 17, dtype=torch.float32).pin_memory().numpy()
        rew = torch.zeros(n_envs, dtype=torch.float32).pin_memory().numpy()
        done = torch.zeros(n_envs, dtype=torch.uint8).pin_memory().numpy()
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            act_np[:] = tape[k]  # this step's inputs reach the pinned buffer inside the timed region
            g.rollout_host(act_np, substeps, obs, rew, done)
        barrier()
        d2h = int(obs.nbytes + rew.nbytes + done.nbytes)
        e2e_how = "tfode_batch_rollout_host: H2D actions, step, D2H obs17 / reward / done, one stream synchronise per step"
    else:
        act_dev = [torch.zeros(n_envs, 2, device="cuda") for _ in range(2)]
        host_rows = [torch.zeros(n_total, abi.ROLLOUT_WIDTH, dtype=torch.float32).pin_memory() for _ in range(2)] if rank == 0 else None
        pend = []
        barrier()
        t0 = time.perf_counter()
        for k in range(K):
            act_np[:] = tape[k]
            act_dev[k & 1].copy_(act, non_blocking=True)
            h = roll.tick_async(substeps, act_dev[k & 1])
            if rank == 0:   # D2H of the gathered rows on the communication stream, behind the collective
                with torch.cuda.stream(roll.comm):
                    host_rows[k & 1].copy_(h.rows, non_blocking=True)
            pend.append(h)
            if len(pend) > 1:
                pend.pop(0).wait()   # at most one gather in flight behind the step being launched
        torch.cuda.synchronize()
        barrier()
        d2h = int(n_total * abi.ROLLOUT_WIDTH * 4)
        e2e_how = ("per step: H2D actions on every rank, step, pack_rollout_kernel, one NCCL all_gather of [n_local, %d] fp32 rows on a second stream "
                   "overlapped with the next step, D2H of all %d rows on the trainer rank" % (abi.ROLLOUT_WIDTH, n_total))
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = units / float(e2e_t.item())

    # -------- CPU baselines beside it (rank 0, N=1 only): bounded sample of the same workload and segment -----------------------------------
    cpu = cpu32 = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cn, dtc, cores, sample = cpu_segment(args.workload, n_envs, substeps, K)
        v = cn * substeps * K / dtc
        cpu = {"value": v, "unit": "env-steps/s", "cores": cores, "per_core": v / cores, "kind": "port", "sample": sample}
        cn, dtc, cores, sample = cpu_segment(args.workload, n_envs, substeps, K, precision="float", budget=CPU_SAMPLE_ENV_STEPS / 2)
        v = cn * substeps * K / dtc
        cpu32 = {"value": v, "unit": "env-steps/s", "cores": cores, "per_core": v / cores, "kind": "port", "sample": sample}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs_sustained", peaks.get("hbm_gbs", 6650.0)))
        kern_ms = tot_ms / K
        traffic = None  # DRAM bytes per launch of the committed ncu capture (same workload, same envs / sub-steps per launch)
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get(args.workload)
            if t and t["envs"] == n_envs and t["substeps"] == substeps:
                traffic = t["dram_bytes_per_launch"]
        except Exception:
            pass
        achieved = ALGO_BYTES_PER_ENV_STEP * n_envs * substeps / (kern_ms * 1e-3) / 1e9

        def seg(lo, hi):  # ms per step averaged over the timed steps that cover episode sub-steps [lo, hi)
            sel = [m for k, m in enumerate(ms) if k * substeps < hi and (k + 1) * substeps > lo]
            return sum(sel) / len(sel) if sel else None
        span = K * substeps
        segments = {"unit": "ms per step of %d sub-steps (rank 0)" % substeps,
                    "first_10pct": seg(0, span // 10), "middle_10pct": seg(span * 4 // 10, span // 2), "last_10pct": seg(span * 9 // 10, span),
                    "mean": sum(ms) / len(ms), "min": min(ms), "max": max(ms)}
        line = {
            "metric": "rover env-steps/sec", "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": tot_ms / K, "higher_is_better": True, "scaling": scaling_of(args.workload), "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_block(args.workload, n_envs, world, substeps, K),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read+write)",
                         "peak_source": ("MEASURED_PEAKS.json " + ("hbm_gbs_sustained" if "hbm_gbs_sustained" in peaks else "hbm_gbs")) if peaks else "fallback 6650",
                         "kernel": "tfode::step_kernel", "kernel_ms": kern_ms,
                         "note": "algorithmic bytes = 1680 B per env-step (SURVEY 8(d)) x envs x sub-steps per launch; one launch per step; the path is bound by "
                                 "shared-memory wavefronts and issue slots of the sequential PGS sweeps, not by HBM (DESIGN.md 4)"},
            "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": int(act_np.nbytes) * world, "d2h_bytes_per_step": d2h, "how": e2e_how},
            "gpu_launches": int(launches), "launches_per_step": launches / K, "clocks": clk, "wall_s_timed_region": t_wall, "segments": segments,
        }
        if parity:
            line["parity"] = parity
        else:
            line["contact_audit"] = {"contact_overflow_steps": int(over), "max_generated_contacts": int(max_gen)}
        if cpu:
            line["cpu_baseline"] = cpu
            line["cpu_baseline_f32"] = cpu32
        if world == 1 and args.workload == "config3" and not args.no_actor_tc:
            line["actor_tc"] = actor_tc_line(peaks)
        print(json.dumps(line))
    roll.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
