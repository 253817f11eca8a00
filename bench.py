#!/usr/bin/env python
"""bench.py — batched ASV env-steps/s on B200 (BASELINE.json config 2), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one pass of the hot path over one batch: every ASV of every env takes one action
(10 physics sub-steps, current field, perception, COLREGs, reward/done — kernel K1), finished
episodes are re-generated (K1b's routine) and re-observed, all in ONE launch
(asvrl_env_step_autoreset; `--unfused` runs the three-launch sequence instead).
  value  = env-steps/s with actions resident in HBM (device time, CUDA events, L2 flushed
           between steps, max over ranks);
  e2e    = the same through BatchedMarineNavEnv.step_host with HOST buffers: pinned actions
           H2D and observations / rewards / status D2H inside the timed region;
  roofline = algorithmic bytes of K1 (280 R + 12 O + 16 C + 8 per env-step)
           / K1's own average launch duration, against MEASURED_PEAKS.json's hbm_gbs;
  cpu_baseline = the float64 C port of the reference (oracle/, pinned against the Python
           reference by tests/golden) on the box's host threads, bounded sample.
           cpu_baseline.learner = the plain-torch restatement of the AC-IQN update on the host.
`--impl reference` times that CPU port alone (the Python reference cannot travel to the GPU
box; DESIGN.md records its speed measured in the build container).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

R, O = 8, 10  # BASELINE config 2: 8 ASVs x 10 buoys per env


def alg_bytes_per_env_step(r, o, c):
    return 280 * r + 12 * o + 16 * c + 8


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.source = index, [], False, "nvidia-smi"

    def _nvml(self):
        """In-process NVML (no subprocess per sample: eight ranks forking nvidia-smi ten times a second
        load the host cores the e2e leg needs); None when pynvml is not importable."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.source = "nvml"
            return nv, h
        except Exception:
            return None

    def run(self):
        nvml = self._nvml()
        while not self.stop_flag:
            try:
                if nvml is not None:
                    nv, h = nvml
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                    act = lambda b: "Active" if bits & b else "Not Active"
                    self.rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)),
                                      str(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)),
                                      act(nv.nvmlClocksEventReasonHwSlowdown),
                                      act(nv.nvmlClocksEventReasonHwThermalSlowdown),
                                      act(nv.nvmlClocksEventReasonSwThermalSlowdown),
                                      act(nv.nvmlClocksEventReasonSwPowerCap)])
                else:
                    r = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                        "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                       timeout=5)
                    if r.returncode == 0 and r.stdout.strip():
                        self.rows.append([c.strip() for c in r.stdout.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        sm = sorted(int(float(r[0])) for r in self.rows if r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows), "source": self.source}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_learner_baseline(seconds=3.0):
    """updates/s of the plain-torch restatement of AGENT:386-432 (oracle/learner_oracle.py, pinned to the
    reference's own train step by tests/test_oracle_golden.py) on torch's host threads, at BASELINE
    config 3 and at the reference's own B=64 / N=8.  A bounded sample of `seconds` per shape."""
    import torch
    from oracle import learner_oracle as lo
    out = {"kind": "port", "cores": torch.get_num_threads(), "dtype": "f32"}
    for name, (B, N) in (("c3_B1024_N32", (1024, 32)), ("ref_B64_N8", (64, 8))):
        ups, n, dt = lo.time_updates(B, N, seconds)
        out[name] = {"updates_per_sec": ups, "sample": f"{n} updates in {dt:.1f} s"}
    return out


def workload_config(E, C, world, stagger=True):
    """The workload both arms run (`config` of the JSON line; identical in the b200 and the reference arm)."""
    return {"workload": f"C2 env-step: {E} envs x {R} ASVs x {O} buoys x {C} vortex cores per GPU; auto-deactivate + auto-reset",
            "envs_per_gpu": E, "robots": R, "buoys": O, "vortex_cores": C, "actions": "continuous U(-1,1)^2",
            "sensor_noise": "Philox4x32-10, in the step", "auto_deactivate": True, "auto_reset": True,
            "episode_clocks": "staggered over [0,1000): ~E/1000 resets per step" if stagger else "synchronised",
            "l2": "GPU arm: flushed between timed steps (256 MiB write) unless --l2 rotate; CPU arm: not applicable",
            "parallelism": f"envs sharded x{world}, no collective"}


class CpuWorkload:
    """The product arm's workload on the float64 C port of the reference (oracle/), all host threads: E envs
    of the config-2 shape generated by the port's reset (ENV:72-176), episode clocks staggered, every step =
    env step with auto-deactivation (TRAINER:168-170) + re-generation of the episodes that ended
    (TRAINER:172, 240-245).  Only the first observation of a fresh episode is left out (the port has no env mask
    for a perception-only pass; ~E/1000 envs per step)."""

    def __init__(self, n_envs, cores, c_cores, seed=0, stagger=True):
        import numpy as np
        from distributional_rl_decision_and_control_b200 import _abi
        from oracle import oracle, scenes
        self.np, self.oracle, self._abi = np, oracle, _abi
        self.E, self.cores, self.C = n_envs, cores, c_cores
        self.p, self.rp = _abi.default_params(), _abi.default_reset_params()
        self.st = oracle.new_state(n_envs, R, O, c_cores)
        self.want = (np.full(n_envs, R, np.int32), np.full(n_envs, c_cores, np.int32), np.full(n_envs, O, np.int32))
        self.min_sg = np.full(n_envs, 30.0)
        self.seed, self.episode, self.step_no = seed + 1, 1, 0
        self._reset(np.ones(n_envs, np.uint8))
        if stagger:
            self.st["ep_steps"][:] = np.random.RandomState(seed).randint(0, 1000, n_envs)
        self.out = oracle.new_out(n_envs, R)
        self.acts = [scenes.random_actions(seed + i, n_envs, R, False) for i in range(4)]
        self.mode = _abi.M_DYNAMICS | _abi.M_AUTO_DEACTIVATE
        self.resets = 0

    def _reset(self, mask):
        self.oracle.env_reset(self.rp, self.st, mask, self.want[0], self.want[1], self.want[2], self.min_sg,
                              self.seed, self.episode, O, self.C)
        self.episode += 1

    def step(self):
        self.step_no += 1
        self.oracle.env_step(self.p, self.st, self.acts[self.step_no % 4], None, self.mode, O, self.C, seed=self.seed,
                             step=self.step_no, n_threads=self.cores, out=self.out, margins=False)
        ended = (self.out["env_status"] & 3) != 0
        n = int(ended.sum())
        if n:
            self.resets += n
            self._reset(ended.astype(self.np.uint8))

    def run(self, n_steps):
        t0 = time.perf_counter()
        for _ in range(n_steps):
            self.step()
        dt = time.perf_counter() - t0
        assert self.np.isfinite(self.out["reward"]).all()
        return self.E * n_steps / dt, dt


def cpu_port_throughput(n_envs, n_steps, cores, c_cores, seed=0, warmup=1):
    """env-steps/s of the float64 C port of the reference on `cores` host threads, on the product arm's
    workload (see CpuWorkload).  Returns (env-steps/s, seconds, resets per step)."""
    w = CpuWorkload(n_envs, cores, c_cores, seed)
    w.run(warmup)
    w.resets = 0
    v, dt = w.run(n_steps)
    return v, dt, w.resets / max(n_steps, 1)


def grad_sync_kind(agent, world):
    if world == 1:
        return "none"
    opt = getattr(agent, "critic_optimizer", None) or getattr(agent, "optimizer", None)
    if getattr(opt, "_peers", None) is not None:
        return ("one-shot all-reduce of the flat bucket over NVLink peer memory inside K7 (pack + flag, "
                "wait + rank-ordered sum + norm, Adam), before clipping")
    return "nccl all-reduce of the flat bucket, before clipping"


def k1_traffic(E, C, fused, key=None):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same command (profiles/k1_traffic.json); None for another shape."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "k1_traffic.json")) as f:
            d = json.load(f)
        if d["envs"] == E and d["cores"] == C and bool(d["fused"]) == bool(fused):
            return d[key] if key else d["dram_bytes_read"] + d["dram_bytes_write"]
    except Exception:
        pass
    return None


def time_learner(agent, K, W, dev, dist, world, eager=False):
    """Learner updates/s: whole Agent update (K2b sample, target + local critic, K3, K4 + backward,
    clip, Adam, actor step) replayed as one CUDA graph when capture works, else eager."""
    import torch
    mode = "graph"
    if eager:
        mode = "eager (--profile)"
    else:
        try:
            agent.enable_cuda_graph()
        except Exception as e:  # e.g. a collective that cannot be captured on this stack
            agent.disable_cuda_graph()
            mode = "eager (%s)" % type(e).__name__
    step = agent.train_graph_replay if agent._graph is not None else (
        agent.train_AC_IQN if agent.agent_type == "AC-IQN" else agent.train_IQN)
    for _ in range(W):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    loss = out[0] if isinstance(out, tuple) else out
    return ms, mode, float(loss)


def time_kernel(fn, K, W, per_graph=20):
    """Device time per launch: `per_graph` launches captured in one CUDA graph and replayed, so the
    figure is the kernel plus its launch gap on the device, not the Python / ctypes call rate."""
    import torch
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(W, 3)):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(per_graph):
            fn()
    reps = max(1, K // per_graph)
    graph.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * per_graph)


def facade_leg(dev, total_timesteps):
    """Trainer.learn through the reference-shaped facade (TRAIN:74-109, TRAINER:85-255) on the example config's
    first curriculum stage; wall clock of the whole loop (reset, per-ASV act, step, replay add, train every 4)."""
    import contextlib
    import io
    import tempfile
    import torch
    from distributional_rl_decision_and_control_b200.agent import Agent
    from distributional_rl_decision_and_control_b200.envs.marinenav.env import MarineNavEnv3
    from distributional_rl_decision_and_control_b200.policy.trainer import Trainer
    sched = {"timesteps": [0], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
    ev = {"num_episodes": [1], "num_robots": [3], "num_cores": [0], "num_obstacles": [0], "min_start_goal_dis": [30.0]}
    with tempfile.TemporaryDirectory() as d, contextlib.redirect_stdout(io.StringIO()):
        train_env = MarineNavEnv3(seed=0, schedule=sched, device=str(dev))
        eval_env = MarineNavEnv3(seed=253, is_eval_env=True, device=str(dev))
        agent = Agent(device=str(dev), seed=100, agent_type="AC-IQN")
        n_train = [0]
        orig = agent.train

        def counted(*a, **k):
            n_train[0] += 1
            return orig(*a, **k)
        agent.train = counted
        tr = Trainer(train_env=train_env, eval_env=eval_env, eval_schedule=ev, rl_agent=agent, imitation=False,
                     il_agent=None)
        tr.save_eval_config(d)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        tr.learn(total_timesteps=total_timesteps, eval_freq=10 ** 9, eval_log_path=d)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    out = {"env_steps_per_sec": train_env.total_timesteps / dt, "updates_per_sec": n_train[0] / dt,
           "env_steps": train_env.total_timesteps, "updates": n_train[0], "seconds": dt,
           "what": "Trainer.learn over MarineNavEnv3 (1 env x 3 ASVs, no buoys) + Agent(AC-IQN, B=64, N=8): the "
                   "unchanged-script path; one kernel launch per env.step, one batch-1 forward per ASV"}
    try:
        with open(os.path.join(ROOT, "profiles", "ref_python_baseline.json")) as f:
            ref = json.load(f)
        c1 = ref["c1_trainer"]
        k = sorted(c1)[0]
        out["python_reference"] = {"env_steps_per_sec": c1[k]["env_steps_per_sec"], "updates_per_sec": c1[k]["updates_per_sec"],
                                   "cores": ref["host"]["cores"], "where": "build container, " + k,
                                   "source": "profiles/ref_python_baseline.json (profiles/probes/ref_python_baseline.py)"}
    except Exception:
        pass
    return out


def training_leg(args, precision, dev, pg, rank, world, barrier, dist, kind="AC-IQN", n_steps=400):
    """BASELINE config 4, per-GPU share: collection policy forward for every ASV (act_batch), the fused
    step, K2a push, and one AC-IQN update (B=1024, N=32) per `update_every` batched steps, all as CUDA
    graphs; the only collective is the gradient all-reduce.  Device-timed, max over ranks."""
    import torch
    from distributional_rl_decision_and_control_b200.agent import Agent
    from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv
    from distributional_rl_decision_and_control_b200.vector_trainer import VectorTrainer
    ET = args.train_envs
    rainbow = kind == "Rainbow"  # its own device-resident prioritised n-step memory, one lane per ASV
    agent = Agent(device=dev, seed=100, agent_type=kind, BATCH_SIZE=1024, num_tau=32, BUFFER_SIZE=1 << 20,
                  process_group=pg, rank=rank, matmul_precision=precision,
                  replay_lanes=ET * R if rainbow else 1)
    env = BatchedMarineNavEnv(ET, R, O, args.cores, device=dev, seed=args.seed + 11, env_offset=rank * ET,
                              fused=True, prefetch="workers")
    vt = VectorTrainer(env, agent, update_every=args.update_every, learning_starts=8, cuda_graph=True)
    vt.learn(40)  # reset, fill the ring, capture both graphs
    update = agent.train_graph_replay if agent._graph is not None else getattr(agent, agent._TRAINERS[kind])
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps0, upd0 = vt.steps, vt.updates
    barrier()
    t0e.record()
    since = 0
    while vt.steps - steps0 < n_steps:
        since += vt.collect(0.05)
        if since >= vt.update_every:
            since = 0
            update()
            vt.updates += 1
    t1e.record()
    barrier()
    tms = torch.tensor([t0e.elapsed_time(t1e)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    tsec = float(tms.item()) * 1e-3
    sync = {"grad_sync": grad_sync_kind(agent, world)}
    if world > 1:
        # driver-side evidence that the exchange kept the replicas bit-identical: every rank's parameter
        # bytes, hashed, all-gathered and compared
        nets = [m for m in (getattr(agent.policy_local, "actor", None), getattr(agent.policy_local, "critic", None),
                            agent.policy_local if isinstance(agent.policy_local, torch.nn.Module) else None) if m is not None]
        flat = torch.cat([p.detach().reshape(-1).view(torch.int32).to(torch.int64) for m in nets for p in m.parameters()])
        w = torch.arange(1, flat.numel() + 1, device=dev, dtype=torch.int64)
        h = torch.stack([flat.sum(), (flat * w).sum()])
        got = [torch.empty_like(h) for _ in range(world)]
        dist.all_gather(got, h)
        sync["replicas_identical"] = bool(all(torch.equal(got[0], x) for x in got))
        for opt in (getattr(agent, "critic_optimizer", None), getattr(agent, "actor_optimizer", None),
                    getattr(agent, "optimizer", None)):
            if hasattr(opt, "check"):
                opt.check()  # raises if any step was skipped (a missed exchange is sticky)
    return {"env_steps_per_sec": world * ET * (vt.steps - steps0) / tsec, **sync,
            "updates_per_sec": (vt.updates - upd0) / tsec, "envs_per_gpu": ET,
            "batched_steps": vt.steps - steps0, "update_every_batched_steps": vt.update_every,
            "batch_per_gpu": 1024, "n_tau": 32, "gemm": precision,
            "mode": "graph" if agent._graph is not None else "eager",
            "what": "act_batch (policy forward for envs x 8 ASVs, eps-greedy) + %s + replay push per "
                    "batched step, one %s update per update_every steps%s"
                    % ("fused step", kind, "; CUDA graphs")}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from distributional_rl_decision_and_control_b200 import _lib
    from distributional_rl_decision_and_control_b200.agent import Agent
    from distributional_rl_decision_and_control_b200.envs.marinenav.batched import BatchedMarineNavEnv

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pinned_cores = None
    if world > 1 and os.environ.get("ASVRL_BENCH_PIN", "1") != "0":
        # N ranks share the host's cores: give every rank its own disjoint slice (the e2e leg is host-launch bound, and an
        # unpinned rank's main thread migrates between cores other ranks' helper threads are using)
        try:
            cores = sorted(os.sched_getaffinity(0))
            per = max(1, len(cores) // world)
            mine = cores[(local * per) % len(cores):(local * per) % len(cores) + per]
            if mine:
                os.sched_setaffinity(0, set(mine))
                pinned_cores = mine
        except (AttributeError, OSError):
            pinned_cores = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    E, C, K, W = args.envs, args.cores, args.steps, max(args.warmup, 3)
    if args.profile:
        K, W, args.no_cpu = min(K, 12), 3, True

    # envs partition across ranks with no data-path collective: rank r owns global envs
    # [r*E, (r+1)*E); the RNG counters carry the global index (results invariant to N)
    # Cold data every step, one of two ways (config.l2 says which):
    #   rotate: NB independent env batches of E envs each are stepped round robin; together their
    #           state, observation, terminal-row and stage buffers exceed the 126 MB L2 (26 MB per
    #           batch at C2), so a batch's data has been evicted when its turn comes again.  Code
    #           stays warm, as it is in any long run.
    #   flush:  one batch, a 256 MiB write between timed steps (also evicts the kernel's code).
    NB = max(1, args.batches) if args.l2 == "rotate" else 1
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    envs = []
    for b in range(NB):
        e_ = BatchedMarineNavEnv(E, R, O, C, device=dev, seed=args.seed + 7919 * b, env_offset=rank * E,
                                 fused=not args.unfused,
                                 prefetch={"stream": "stream", "workers": "workers", "off": False}[args.prefetch])
        e_.reset()
        if not args.no_stagger:
            # steady state: episode clocks spread over [0, 1000) so ~E/1000 episodes time out per step,
            # as in a long run; without it every env of the batch times out in the same step
            e_.state["ep_steps"].copy_(torch.randint(0, 1000, (E,), device=dev, generator=gen, dtype=torch.int32))
        envs.append(e_)
    env = envs[0]
    pool = [torch.rand((E, R, 2), device=dev, generator=gen) * 2 - 1 for _ in range(8)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if NB == 1 else None  # > 126 MB L2

    def cold():
        if flush is not None:
            flush.zero_()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- (1) eager pass, K1 timed alone (the roofline numerator) ----
    for i in range(W * NB):
        envs[i % NB].step(pool[i % 8])
    KE = min(K, 400)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(KE)]
    barrier()
    for i in range(KE):
        cold()
        en = envs[i % NB]
        if en.fused:
            ev[i][0].record()
            en.step_fused(pool[i % 8])     # K1 with the fused auto-reset tail: the whole step
            ev[i][1].record()
        else:
            ev[i][0].record()
            en.step_no_reset(pool[i % 8])  # K1
            ev[i][1].record()
            en.finish_episodes()           # K1b + masked perception-only K1
    barrier()
    k1_all = sorted(e[0].elapsed_time(e[1]) for e in ev)
    k1_ms = k1_all[len(k1_all) // 2]  # median launch: robust at the driver's --steps 20 (one cold outlier is 5 % of a mean)

    # ---- (2) the product path: one CUDA graph per step (the fused launch, or K1, K1b, masked K1) ----
    # the fused step is a single launch, issued directly; the three-launch sequence is one graph
    use_graph = not env.fused
    if use_graph:
        for en in envs:
            en.enable_cuda_graph(True)
        do_step = lambda en, a: en.step_graph(a)
    else:
        do_step = lambda en, a: en.step(a)
    for i in range(W * NB):
        do_step(envs[i % NB], pool[i % 8])
    sampler = ClockSampler(local)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(K)]
    launches0 = sum(en.launches for en in envs)
    barrier()
    sampler.start()
    for i in range(K):
        cold()  # flush mode: L2 flush between timed iterations (not timed)
        ev[i][0].record()
        do_step(envs[i % NB], pool[i % 8])
        ev[i][1].record()
    if getattr(env, "prefetch", None) == "stream":  # staging still in flight belongs to the timed work
        tail0, tail1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tail0.record()
        for en in envs:
            torch.cuda.current_stream().wait_stream(en._side)
        tail1.record()
    barrier()
    per_step = sorted(e[0].elapsed_time(e[1]) for e in ev)
    step_sum_ms = sum(per_step)
    if getattr(env, "prefetch", None) == "stream":
        step_sum_ms += tail0.elapsed_time(tail1)
    # value: this rank's MEDIAN step time x K (then the max over ranks); the plain sum of the K steps is kept
    # beside it (`ms_per_step_mean`) — at K = 20 one slow launch on one rank otherwise sets the whole-job figure
    step_ms = per_step[len(per_step) // 2] * K
    launches = sum(en.launches for en in envs) - launches0

    # cross-check without the flush gaps: K steps back to back, the side stream joined at the end
    b2b0, b2b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    b2b0.record()
    for i in range(K):  # one batch only: everything L2-resident (NOT the reported value)
        env.step(pool[i % 8]) if env.fused else env.step_graph(pool[i % 8])
    if getattr(env, "prefetch", None) == "stream":
        torch.cuda.current_stream().wait_stream(env._side)
    b2b1.record()
    barrier()
    b2b_ms = b2b0.elapsed_time(b2b1) / K

    # ---- (3) e2e: host buffers in, host buffers out, through the public API ----
    # Two shards of E/2 envs (send / recv): the D2H copy of one overlaps the kernel of the other.
    # Every step of every shard: pinned actions H2D, the fused launch, observations + rewards +
    # status D2H, and the caller touches the results (a checksum of the rewards) before it sends
    # the shard's next actions.
    from distributional_rl_decision_and_control_b200.envs.marinenav.host_pipeline import HostPipelinedEnv
    G = args.e2e_groups
    henv = HostPipelinedEnv(E, R, O, C, groups=G, device=dev, seed=args.seed + 3, env_offset=rank * E,
                            fused=not args.unfused)
    if not args.no_stagger:
        for sh in henv.shards:
            sh.state["ep_steps"].copy_(torch.randint(0, 1000, (sh.E,), device=dev, generator=gen, dtype=torch.int32))
    host_actions = [[p[g * henv.E:(g + 1) * henv.E].cpu().pin_memory() for g in range(G)] for p in pool]
    if not args.unfused and not args.e2e_eager:
        henv.enable_cuda_graph(zero_copy=args.e2e_zero_copy)  # one graph launch per send: [actions H2D, fused step, results D2H]
    torch.cuda.synchronize()
    for i in range(W):
        for g in range(G):
            henv.send(g, host_actions[i % 8][g])
    for g in range(G):
        henv.recv(g)
    barrier()
    reward_np = [henv.host_out[g]["reward"].numpy() for g in range(G)]
    t0 = time.perf_counter()
    for g in range(G):
        henv.send(g, host_actions[0][g])
    checksum = 0.0
    for i in range(1, K + 1):
        for g in range(G):
            henv.recv(g)                            # results of shard g's step i-1 are on the host
            checksum += float(reward_np[g][0, 0])   # the caller reads them (numpy view of the pinned block)
            if i < K:
                henv.send(g, host_actions[i % 8][g])
    e2e_s = time.perf_counter() - t0
    barrier()
    e2e_d2h = G * henv.d2h_bytes
    finite = bool(torch.isfinite(env.out["reward"]).all().item())

    # ---- (4) learner leg (BASELINE config 3 and the reference's own B=64 / N=8) ----
    pg = dist.group.WORLD if world > 1 else None
    learner = {}
    big = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=1024, num_tau=32,
                BUFFER_SIZE=1 << 20, process_group=pg, rank=rank)
    env2 = BatchedMarineNavEnv(E, R, O, C, device=dev, seed=args.seed + 1, env_offset=rank * E, fused=False)
    env2.reset()
    for i in range(24):  # fill the replay shard from real transitions (K2a)
        obs = env2.obs
        out = env2.step_no_reset(pool[i % 8])
        big.memory.push_step(obs, pool[i % 8], False, out)
        env2.finish_episodes()
    small = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8,
                  BUFFER_SIZE=1 << 20, process_group=pg, rank=rank)
    small.memory.ring = big.memory.ring  # same shard, two batch shapes
    fast = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=1024, num_tau=32,
                 BUFFER_SIZE=1 << 20, process_group=pg, rank=rank, matmul_precision="tf32")
    fast.memory.ring = big.memory.ring
    KL = 3 if args.profile else max(50, min(K, 500))
    for name, ag in (("c3_B1024_N32", big), ("c3_B1024_N32_tf32", fast), ("ref_B64_N8", small)):
        ms, mode, loss = time_learner(ag, KL, 2 if args.profile else 10, dev, dist, world, eager=args.profile)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        learner[name] = {"updates_per_sec": KL / (ms * 1e-3), "ms_per_update": ms / KL, "mode": mode,
                         "batch_per_gpu": ag.BATCH_SIZE, "n_tau": ag.num_tau, "last_critic_loss": loss,
                         "gemm": ag.matmul_precision,
                         "grad_sync": grad_sync_kind(ag, world)}
    # K3 / K4 alone at C3 (their own rooflines)
    B3, N3 = 1024, 32
    g = torch.Generator(device=dev).manual_seed(5)
    z, zn = torch.randn((B3, N3), device=dev, generator=g), torch.randn((B3, N3), device=dev, generator=g)
    tau = torch.rand((B3, N3), device=dev, generator=g)
    rr, dd = torch.randn((B3,), device=dev, generator=g), torch.zeros((B3,), device=dev)
    ws = torch.zeros(1, dtype=torch.int32, device=dev)
    o_l, o_g, o_s = torch.empty(B3, device=dev), torch.empty((B3, N3), device=dev), torch.empty(1, device=dev)
    k4_ms = time_kernel(lambda: _lib.qhuber_fwd_bwd(z, tau, zn, rr, dd, 0.99, 1.0, ws, o_l, o_g, o_s), 20 if args.profile else 200, 20)
    t_o, c_o = torch.empty((B3, N3), device=dev), torch.empty((B3, N3, 64), device=dev)
    k3_ms = time_kernel(lambda: _lib.iqn_tau_cos(B3, N3, 64, dev, seed=1, counter=1, tau_out=t_o, cos_out=c_o), 20 if args.profile else 200, 20)
    peak, which = measured_peak()
    k4_bytes = 4 * B3 * (2 * N3 + N3 + 2) + 4 * B3 * (N3 + 1)
    k3_bytes = 4 * B3 * N3 * (1 + 64)
    learner["k4_qhuber"] = {"ms": k4_ms, "algorithmic_bytes": k4_bytes, "achieved_gbs": k4_bytes / (k4_ms * 1e-3) / 1e9,
                            "frac": k4_bytes / (k4_ms * 1e-3) / 1e9 / peak,
                            "note": "graph of 20 back-to-back launches; 0.5 MB per launch is latency bound"}
    learner["k3_tau_cos"] = {"ms": k3_ms, "algorithmic_bytes": k3_bytes, "achieved_gbs": k3_bytes / (k3_ms * 1e-3) / 1e9,
                             "frac": k3_bytes / (k3_ms * 1e-3) / 1e9 / peak}
    # K8 alone: one dense layer of the collection forward (65 536 ASVs, 256 -> 128, bias + ReLU) on tcgen05
    # (3 x TF32 split) against the cuBLAS sgemm + bias/ReLU passes it replaces
    Mk = 65536
    xk = torch.randn((Mk, 256), device=dev, generator=g)
    wk = torch.randn((128, 256), device=dev, generator=g) / 16
    bk = torch.randn((128,), device=dev, generator=g)
    ok = torch.empty((Mk, 128), device=dev)
    k8_ms = time_kernel(lambda: _lib.linear_tf32x3(xk, wk, bk, relu=True, out=ok), max(K, 200), W)
    sg_ms = time_kernel(lambda: torch.relu_(torch.addmm(bk, xk, wk.t(), out=ok)), max(K, 200), W)
    k8_bytes = 4 * Mk * (256 + 128)
    learner["k8_linear_tf32x3"] = {"ms": k8_ms, "rows": Mk, "in_dim": 256, "out_dim": 128,
                                   "algorithmic_bytes": k8_bytes, "achieved_gbs": k8_bytes / (k8_ms * 1e-3) / 1e9,
                                   "frac": k8_bytes / (k8_ms * 1e-3) / 1e9 / peak,
                                   "tflops_fp32_equivalent": 2 * Mk * 256 * 128 / (k8_ms * 1e-3) / 1e12,
                                   "cublas_sgemm_bias_relu_ms": sg_ms}
    # K9 alone: the fused collection policy forward for 65 536 ASVs (the per-GPU share of config 4) against the
    # K6 + 2 x K8 + actor-head sequence of round 1; algorithmic HBM bytes 148 in + 8 out per ASV
    try:
        from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import Actor as _Actor
        kag = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
        Mp = 65536
        xo = (torch.randn((Mp, 7), device=dev, generator=g), torch.randn((Mp, 5, 5), device=dev, generator=g),
              (torch.rand((Mp, 5), device=dev, generator=g) < 0.6).float())
        k9_ms = time_kernel(lambda: kag.act_batch(xo, eps=0.05), max(K, 200), W)
        _lib.policy_forward_status()
        _Actor.fused_policy = False
        old_ms = time_kernel(lambda: kag.act_batch(xo, eps=0.05), max(K, 200), W)
        _Actor.fused_policy = True
        k9_bytes = Mp * (148 + 8)
        k9_flop = 2.0 * Mp * (56 * 7 + 200 * 5 + 256 * 128 + 128 * 128 + 2 * 128)
        learner["k9_policy_forward"] = {
            "ms": k9_ms, "rows": Mp, "algorithmic_bytes": k9_bytes, "achieved_gbs": k9_bytes / (k9_ms * 1e-3) / 1e9,
            "frac_hbm": k9_bytes / (k9_ms * 1e-3) / 1e9 / peak, "tflops_fp32_equivalent": k9_flop / (k9_ms * 1e-3) / 1e12,
            "tensor_tflops_issued": 3 * 2.0 * Mp * (256 * 128 + 128 * 128) / (k9_ms * 1e-3) / 1e12,
            "round1_k6_k8_head_ms": old_ms, "launches": 2,
            "note": "one persistent tcgen05 launch (+ the weight-pack launch): fp16 two-piece split, 3 products"}
    except Exception as e:
        learner["k9_policy_forward"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    # K10 / K11 / K12 alone at C3: the critic trunk forward, backward chain and weight gradients on tcgen05 against the K3 +
    # cuBLAS sgemm + K5 sequence they replace.  Algorithmic HBM bytes per quantile row: the operand images (fp16 hi | lo)
    # — cos 256, H1 1024, H2 512, R3 512 saved by the differentiable forward; dP3 512, dP2 512, dP1 1024, aux 256 written
    # by K11; K12 reads all eight.
    try:
        from distributional_rl_decision_and_control_b200.policy.iqn_ops import critic_trunk as _old_trunk
        cag = Agent(device=dev, seed=100, agent_type="AC-IQN", BATCH_SIZE=64, num_tau=8, BUFFER_SIZE=256)
        cr = cag.policy_local.critic
        cps = (cr.cos_embedding.weight, cr.cos_embedding.bias, cr.hidden_layer.weight, cr.hidden_layer.bias,
               cr.hidden_layer_2.weight, cr.hidden_layer_2.bias, cr.output_layer.weight.reshape(-1), cr.output_layer.bias)
        Bc, Nc = 1024, 32
        Mc = Bc * Nc
        tau_c = torch.rand((Bc, Nc), device=dev, generator=g)
        feats_c = torch.relu(torch.randn((Bc, 256), device=dev, generator=g) + 0.4).contiguous()
        aenc_c = torch.relu(torch.randn((Bc, 128), device=dev, generator=g) + 0.4).contiguous()
        dz_c = (torch.randn((Bc, Nc), device=dev, generator=g) / Mc).contiguous()
        cws = _lib.critic_workspace("bench", dev)
        reps = 20 if args.profile else 200
        k10_ms = time_kernel(lambda: _lib.critic_forward(feats_c, aenc_c, *cps, Nc, cws, tau_in=tau_c), reps, 20)
        k10s_ms = time_kernel(lambda: _lib.critic_forward(feats_c, aenc_c, *cps, Nc, cws, tau_in=tau_c, save=True), reps, 20)
        _, _, (ci, h1i, h2i, r3i) = _lib.critic_forward(feats_c, aenc_c, *cps, Nc, cws, tau_in=tau_c, save=True)
        fmx = feats_c.amax().reshape(1)
        k11_ms = time_kernel(lambda: _lib.critic_backward(dz_c, r3i, h2i, h1i, feats_c, aenc_c, cps[2], cps[4], cps[6], Nc, fmax=fmx), reps, 20)
        k11a_ms = time_kernel(lambda: _lib.critic_backward(dz_c, r3i, h2i, None, feats_c, aenc_c, cps[2], cps[4], cps[6], Nc, full=False), reps, 20)
        _, _, (d3, d2, d1, ax), sc = _lib.critic_backward(dz_c, r3i, h2i, h1i, feats_c, aenc_c, cps[2], cps[4], cps[6], Nc, fmax=fmx)
        k12_ms = time_kernel(lambda: _lib.critic_wgrad(d3, d2, d1, ax, ci, h1i, h2i, r3i, sc, Bc, Nc), reps, 20)
        torch.cuda.synchronize()
        _lib.critic_forward_status(); _lib.critic_backward_status(); _lib.critic_wgrad_status()

        def _old():
            with torch.no_grad():
                _, cs = _lib.iqn_tau_cos(Bc, Nc, 64, dev, tau_in=tau_c)
                return _old_trunk(cs.reshape(Mc, 64), feats_c, aenc_c, cps[0], cps[1], cps[2], cps[3], cps[4], cps[5],
                                  cr.output_layer.weight, cps[7], Nc)
        old_ms = time_kernel(_old, reps, 20)
        flop = 2.0 * Mc * (64 * 256 + 256 * 128 + 128 * 128 + 128)
        gbs = lambda nbytes, ms: nbytes / (ms * 1e-3) / 1e9
        b_fwd, b_save = Mc * 8 + Bc * 1536, Mc * (8 + 2304) + Bc * 1536
        b_k11, b_k12 = Mc * (2048 + 2304 + 4) + Bc * 3072, Mc * 4608
        learner["k10_critic_forward"] = {
            "ms": k10_ms, "ms_save": k10s_ms, "rows": Mc, "tflops_fp32_equivalent": flop / (k10_ms * 1e-3) / 1e12,
            "tensor_tflops_issued": 3 * flop / (k10_ms * 1e-3) / 1e12, "algorithmic_bytes": b_fwd,
            "algorithmic_bytes_save": b_save, "achieved_gbs_save": gbs(b_save, k10s_ms), "frac_hbm_save": gbs(b_save, k10s_ms) / peak,
            "k3_sgemm_k5_ms": old_ms,
            "note": "bound: tensor issue + CUDA-core epilogue (the inference pass moves 2 MB); save = + the operand images"}
        learner["k11_critic_backward"] = {
            "ms": k11_ms, "ms_actor_pass": k11a_ms, "rows": Mc, "algorithmic_bytes": b_k11, "achieved_gbs": gbs(b_k11, k11_ms),
            "frac_hbm": gbs(b_k11, k11_ms) / peak, "tflops_fp32_equivalent": 2.0 * Mc * (128 * 128 + 128 * 256) / (k11_ms * 1e-3) / 1e12}
        learner["k12_critic_wgrad"] = {
            "ms": k12_ms, "rows": Mc, "algorithmic_bytes": b_k12, "achieved_gbs": gbs(b_k12, k12_ms),
            "frac_hbm": gbs(b_k12, k12_ms) / peak, "bound": "hbm (operand images read once)",
            "tflops_fp32_equivalent": 2.0 * Mc * (128 * 128 + 128 * 256 + 256 * 64 + 3 * 128) / (k12_ms * 1e-3) / 1e12}
    except Exception as e:
        learner["k10_critic_forward"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    # ---- facade leg: the path train_RL_agents.py drives UNCHANGED — one MarineNavEnv3 (config-1 shape: 3 ASVs,
    # no buoys) + Agent (B=64, N=8) + Trainer.learn — timed as a whole loop, next to the Python reference's
    # own figures for the same loop (profiles/ref_python_baseline.json, produced in the build container)
    facade = None
    if not args.profile and rank == 0:
        try:
            facade = facade_leg(dev, args.facade_steps)
        except Exception as e:
            facade = {"error": f"{type(e).__name__}: {e}"[:300]}
    # ---- (5) training leg (BASELINE config 4, per-GPU share): collection policy forward for every ASV,
    # the fused step, K2a push, and one AC-IQN update (B=1024, N=32) per `update_every` batched steps,
    # all as CUDA graphs; the only collective is the gradient all-reduce.  Device-timed.
    training = None
    if not args.profile:
        training = {}
        for prec in ("fp32", "tf32"):
            try:
                training[prec] = training_leg(args, prec, dev, pg, rank, world, barrier, dist)
            except Exception as e:  # the headline legs above must not be lost to this one
                training[prec] = {"error": f"{type(e).__name__}: {e}"[:300]}
    # ---- (6) BASELINE config 5: the same loop for the other five agents of the reference, one after
    # the other (fp32, every world size: the six-agent 1/2/4/8 study; a leg that fails is reported, not fatal)
    agents = None
    if not args.profile and not args.no_agents:
        agents = {}
        for kind in ("IQN", "DDPG", "DQN", "SAC", "Rainbow"):
            try:
                agents[kind] = training_leg(args, "fp32", dev, pg, rank, world, barrier, dist, kind=kind, n_steps=200)
            except Exception as e:
                agents[kind] = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    times = torch.tensor([step_ms, k1_ms, e2e_s * 1e3, step_sum_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    step_ms, k1_ms, e2e_ms, step_sum_ms = (float(x) for x in times.tolist())

    if rank == 0:
        # issue-rate roof of the dominant kernel (it is issue bound, not HBM bound): warp instructions
        # of one launch (ncu, profiles/k1_traffic.json) / (SMs x 4 schedulers x SM clock)
        clk = sampler.summary()
        winst = k1_traffic(E, C, env.fused, key="warp_instructions")
        issue = None
        if winst and clk.get("sm_mhz"):
            sms = torch.cuda.get_device_properties(dev).multi_processor_count
            floor_ms = winst / (sms * 4 * clk["sm_mhz"] * 1e6) * 1e3
            issue = {"warp_instructions": winst, "issue_slots_per_s": sms * 4 * clk["sm_mhz"] * 1e6,
                     "floor_ms": floor_ms, "frac": floor_ms / k1_ms}
        per_launch = alg_bytes_per_env_step(R, O, C) * E
        achieved = per_launch / (k1_ms * 1e-3) / 1e9
        def _num(d, *keys):
            for k in keys:
                d = d.get(k) if isinstance(d, dict) else None
            return round(d, 6) if isinstance(d, float) else d
        tr32 = (training or {}).get("fp32") or {}
        # the driver's record keeps `config`, `roofline`, `e2e`, `cpu_baseline`, `clocks` and drops other top-level
        # keys: the numbers the review needs are mirrored here in compact form
        also = {
            "ms_per_step_mean": step_sum_ms / K, "l2_resident_ms_per_step": b2b_ms,
            "asv_steps_per_sec": world * E * R * K / (step_ms * 1e-3),
            "training_fp32_env_steps_per_sec": _num(training, "fp32", "env_steps_per_sec"),
            "training_tf32_env_steps_per_sec": _num(training, "tf32", "env_steps_per_sec"),
            "training_fp32_updates_per_sec": _num(training, "fp32", "updates_per_sec"),
            "grad_sync": tr32.get("grad_sync"), "replicas_identical": tr32.get("replicas_identical"),
            "k10_critic_forward_ms": _num(learner, "k10_critic_forward", "ms"),
            "k10_critic_forward_save_ms": _num(learner, "k10_critic_forward", "ms_save"),
            "k10_replaces_k3_sgemm_k5_ms": _num(learner, "k10_critic_forward", "k3_sgemm_k5_ms"),
            "k11_critic_backward_ms": _num(learner, "k11_critic_backward", "ms"),
            "k11_frac_hbm": _num(learner, "k11_critic_backward", "frac_hbm"),
            "k12_critic_wgrad_ms": _num(learner, "k12_critic_wgrad", "ms"),
            "k12_frac_hbm": _num(learner, "k12_critic_wgrad", "frac_hbm"),
            "learner_c3_fp32_ms": _num(learner, "c3_B1024_N32", "ms_per_update"),
            "learner_c3_tf32_ms": _num(learner, "c3_B1024_N32_tf32", "ms_per_update"),
            "learner_ref_B64_N8_ms": _num(learner, "ref_B64_N8", "ms_per_update"),
            "k9_policy_forward_ms": _num(learner, "k9_policy_forward", "ms"),
            "k9_frac_hbm": _num(learner, "k9_policy_forward", "frac_hbm"),
            "k9_tflops_fp32_equivalent": _num(learner, "k9_policy_forward", "tflops_fp32_equivalent"),
            "k4_qhuber_ms": _num(learner, "k4_qhuber", "ms"), "k3_tau_cos_ms": _num(learner, "k3_tau_cos", "ms"),
            "agents_c5_env_steps_per_sec": {k: _num(v, "env_steps_per_sec") for k, v in (agents or {}).items()},
            "facade_env_steps_per_sec": _num(facade, "env_steps_per_sec"),
            "facade_updates_per_sec": _num(facade, "updates_per_sec"),
            "facade_python_reference_env_steps_per_sec": _num(facade, "python_reference", "env_steps_per_sec"),
        }
        line = {
            "metric": "env_steps_per_sec", "value": world * E * K / (step_ms * 1e-3), "unit": "env-steps/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": step_ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": workload_config(E, C, world, not args.no_stagger),
            "impl_notes": {"launch": ("one fused launch per step (K1 + re-generation + first observation), no graph"
                                      if env.fused else "one CUDA graph per step (K1, K1b, masked K1)"),
                           "next_episodes": (f"pre-generated on a side stream every {env.prefetch_every} steps "
                                             "(asvrl_env_stage_refill)" if getattr(env, "prefetch", None) == "stream"
                                             else "pre-generated by refill CTAs of the step launch"
                                             if getattr(env, "prefetch", None) == "workers" else "generated in place"),
                           "value": "median step time of this rank x K, max over ranks (ms_per_step_mean: the plain mean)",
                           "l2": (f"inputs larger than L2: {NB} env batches of {E} envs stepped round robin, "
                                  f"~{26 * NB} MB of state / observation / stage buffers" if NB > 1
                                  else "flushed between timed steps (256 MiB write)")},
            "e2e": {"value": world * E * K / (e2e_ms * 1e-3), "unit": "env-steps/s",
                    "h2d_bytes_per_step": E * R * 2 * 4, "d2h_bytes_per_step": e2e_d2h,
                    "api": ("HostPipelinedEnv.send/recv, %d shards of E/%d envs, one CUDA graph launch per send "
                           "(actions H2D, fused step, results D2H); obs_self + obs_objects + reward + "
                           "status per ASV (the mask is objects[...,4] != 0, as in the reference's step())") % (G, G),
                    "host_cores_per_rank": pinned_cores if pinned_cores else "unpinned",
                    "d2h_gbs_per_rank": e2e_d2h / (e2e_ms * 1e-3 / K) / 1e9,
                    "limiter": "PCIe D2H of the reference-shaped float32 outputs (%.2f MB per step and rank): one rank alone "
                               "copies at 48 GB/s (%.0f us per step); with 8 ranks copying at once the host fabric gives "
                               "12.5 GB/s to ranks 0-3 and 21 GB/s to ranks 4-7 (134 GB/s aggregate = at most 123 M "
                               "env-steps/s over the box; profiles/r02_pcie_scaling_n8.txt)"
                               % (e2e_d2h / 1e6, e2e_d2h / 48e9 * 1e6)},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": k1_traffic(E, C, env.fused), "peak_source": which,
                         "kernel": "env_step_kernel<false,true,1> (fused auto-reset, config-2 shape instantiation)" if env.fused
                                   else "env_step_kernel<false,false>", "kernel_ms": k1_ms,
                         "kernel_ms_is": "median of the event-timed launches",
                         "algorithmic_bytes_per_launch": per_launch,
                         "issue": issue,
                         "note": "instruction-issue bound by design (SURVEY.md F5): issue.frac = share of issue slots used. "
                                 "The launch's 6.7 MB of stores stay dirty in L2 and are written back during the "
                                 "UN-timed flush between steps: the timed region never pays them (traffic = reads only)",
                         "also": also},
            "learner": learner, "training": training, "agents_c5": agents, "facade": facade,
            "clocks": sampler.summary(), "finite": finite,
        }
        if world == 1 and not args.no_cpu:
            cores = host_threads()
            v, dt, rps = cpu_port_throughput(E, args.cpu_steps, cores, C)
            line["cpu_baseline"] = {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                                    "sample": f"{E} envs x {args.cpu_steps} steps of the same workload ({rps:.1f} resets "
                                              f"per step), float64 C port of the reference, {dt:.1f} s",
                                    "learner": cpu_learner_baseline()}
        print(json.dumps(line), flush=True)
    if world > 1:
        # captured graphs hold NCCL work: tearing the communicator down under them can hang, so
        # every rank meets once more and leaves without the destructor chain
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def run_reference(args):
    """The reference's CPU implementation of the path, as its float64 C port (the Python reference does not
    travel to the GPU box; its own speed, measured in the build container, is profiles/ref_python_baseline.json),
    on all host threads, on the SAME workload as the b200 arm (CpuWorkload).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = host_threads()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    K, W = args.steps, max(args.warmup, 1)
    E = args.envs
    v, dt, rps = cpu_port_throughput(E, K, cores, args.cores, warmup=W)
    line = {"impl": "reference", "metric": "env_steps_per_sec", "value": v, "unit": "env-steps/s",
            "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(E, args.cores, world, True),
            "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                             "sample": f"{E} envs x {K} steps ({rps:.1f} resets per step), float64 C port of the reference",
                             "learner": cpu_learner_baseline()},
            "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU (BASELINE config 2: 4096)")
    ap.add_argument("--cores", type=int, default=8, help="vortex cores per env (the current disturbance)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--cpu-steps", type=int, default=1000, help="steps of the cpu_baseline sample (same envs as --envs)")
    ap.add_argument("--facade-steps", type=int, default=3000, help="Trainer.learn budget of the facade leg (learning starts at 2000)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--l2", default="flush", choices=["rotate", "flush"],
                    help="how every timed step sees cold data: rotate env batches (> L2) or flush L2")
    ap.add_argument("--batches", type=int, default=8, help="env batches rotated with --l2 rotate")
    ap.add_argument("--prefetch", default="stream", choices=["stream", "workers", "off"])
    ap.add_argument("--no-stagger", action="store_true", help="start every episode clock at 0")
    ap.add_argument("--unfused", action="store_true", help="three launches per step (K1, K1b, masked K1)")
    ap.add_argument("--train-envs", type=int, default=8192, help="envs per GPU of the training leg (config 4: 65536 / 8)")
    ap.add_argument("--update-every", type=int, default=4, help="batched steps per learner update in the training leg")
    ap.add_argument("--no-agents", action="store_true", help="skip the config-5 leg (the other five agents)")
    ap.add_argument("--e2e-groups", type=int, default=4, help="shards of the host pipeline (send / recv)")
    ap.add_argument("--e2e-zero-copy", action="store_true",
                    help="the step kernel stores its results straight into pinned host memory (no D2H copy node)")
    ap.add_argument("--e2e-eager", action="store_true", help="e2e leg without the per-shard CUDA graphs")
    ap.add_argument("--profile", action="store_true",
                    help="short run for ncu: few env steps, three EAGER learner updates per shape (no graph), "
                         "no e2e / CPU legs; the numbers it prints are not bench values")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
