#!/usr/bin/env python
"""bench.py — env-steps/s of the fused step pipeline (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload NAME]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one control step of the path over one batch of N_env synthetic environments:
dex_pre_physics + dex_post_physics (with the fused reset of finished envs and the step epilogue fused into its last CTA),
physics replaced by nothing (PhysX is out of scope; the sim-state tensors are static synthetic
data already resident in HBM).  Default workload = BASELINE.json configs[3]: BlindGrasping,
16 384 envs per GPU (the config the 1/2/4/8-GPU metric is quoted on); envs shard across ranks with
no collective on the step path ("weak" scaling) and ONE NCCL all-reduce of the statistics vector
per timed region.

L2 protocol: the timed loop rotates over a ring of R independent env replicas (each with its own
sim tensors, path state and outputs) whose total footprint exceeds 2x the 126 MB L2, so every step
streams its inputs from HBM ("inputs larger than L2").

The CPU oracle (oracle/) is imported ONLY by the cpu_baseline leg and by --impl reference.
"""
from __future__ import annotations

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

import torch  # noqa: E402

METRIC = "env_steps_per_s_fused_step_pipeline"
UNIT = "env-steps/s"
L2_BYTES = 126 * 1024 * 1024

# name -> (task, envs per GPU, algorithmic bytes per env-step: pre, post)   (SURVEY §8d / BASELINE.md §3)
WORKLOADS = {
    "blindgrasping_16384": ("BlindGrasping", 16384, 464, 1997),
    "blindgrasping_4096": ("BlindGrasping", 4096, 464, 1997),
    "basetask_4096": ("BaseTask", 4096, 464, 2455),
    "basetask_64": ("BaseTask", 64, 464, 2455),
}
DEFAULT_WORKLOAD = "blindgrasping_16384"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--envs", type=int, default=0, help="override envs per GPU")
    ap.add_argument("--observer", type=int, default=0, help="1 = fused RewardComponentObserver on (SURVEY §8f N1)")
    ap.add_argument("--nb-hand", type=int, default=31, help="hand bodies per env (31 = survey layout; 37 = sensitivity line)")
    ap.add_argument("--obs-norm", type=int, default=0, help="1 = fused policy-input normalisation on (SURVEY §8f N3)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tile-kernel", action="store_true",
                    help="A/B: run the round-1 one-tile-per-CTA kernel (DEX_FLAG_TILE_KERNEL) instead of the persistent v2 kernel")
    ap.add_argument("--single-pass", action="store_true",
                    help="DEX_FLAG_SINGLE_PASS on every post-physics launch: the single-pass kernel where it applies (2-4 tiles per SM)")
    ap.add_argument("--sweep", action="store_true", help="also print the post-physics num_envs sweep (configs[4])")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the bounded CPU-baseline sample")
    return ap.parse_args()


def measured_peak_hbm():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while a timed region runs."""

    def __init__(self, index: int, period_s: float = 0.01):
        self.index, self.period, self.samples, self._stop = index, period_s, [], threading.Event()
        self.sm_max, self._thread, self._nvml = None, None, None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _read(self):
        n = self._nvml
        mhz = int(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM))
        try:
            reasons = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._h))
        except Exception:
            reasons = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
        return mhz, reasons

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self._read())
            except Exception:
                pass
            self._stop.wait(self.period)

    def start(self):
        if self._nvml is not None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
            self._thread = None

    def summary(self):
        if self._nvml is None:
            return _clocks_from_nvidia_smi(self.index)
        names = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
                 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
                 0x100: "display_clock_setting"}
        if not self.samples:
            try:
                self.samples.append(self._read())
            except Exception:
                return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        mhz = sorted(s[0] for s in self.samples)
        bits = 0
        for _, r in self.samples:
            bits |= r
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": self.sm_max,
                "reasons": [v for k, v in names.items() if bits & k], "samples": len(self.samples)}


def _clocks_from_nvidia_smi(index):
    try:
        out = subprocess.run(["nvidia-smi", f"--id={index}", "--query-gpu=clocks.sm,clocks.max.sm,"
                              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
        f = [x.strip() for x in out.strip().split(",")]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        return {"sm_mhz": int(f[0]), "sm_max_mhz": int(f[1]),
                "reasons": [n for n, v in zip(names, f[2:]) if v.lower().startswith("active")], "samples": 1}
    except Exception:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}


# ------------------------------------------------------------------------------------------------
# product arm: ring of env replicas on one GPU
# ------------------------------------------------------------------------------------------------
class Ring:
    """R independent replicas of the workload (own sim tensors, path state, outputs, actions)."""

    def __init__(self, task, n, device, seed, observer, min_bytes=2 * L2_BYTES, max_replicas=64, obs_norm=False,
                 nb_hand=31):
        from dexrobot_isaac_b200 import synthetic
        from dexrobot_isaac_b200.config import load_task_config
        from dexrobot_isaac_b200.env import DexStepEnv
        from dexrobot_isaac_b200.params import BodyLayout

        self.task, self.n, self.device = task, n, device
        self.layout = BodyLayout.default(with_box=task == "BlindGrasping", nb_hand=nb_hand)
        self.cfg = load_task_config(task, num_envs=n)
        gen = torch.Generator(device=device).manual_seed(seed)
        self.envs, self.actions = [], []

        def add():
            env = DexStepEnv(self.cfg, n, self.layout, synthetic.dof_props(), device=device, observer=bool(observer),
                             seed=seed + len(self.envs), sim=synthetic.random_state(n, self.layout, gen, device))
            env.pipeline.uniforms = None  # resets draw from the in-kernel Philox stream
            if obs_norm:
                d = env.pipeline.spec.obs_dim
                env.set_obs_normalization(torch.zeros(d, dtype=torch.float64), torch.ones(d, dtype=torch.float64))
            env.reset()
            a = synthetic.random_actions(n, env.num_actions, gen, device).contiguous()
            for _ in range(2):
                env.pipeline.step(a)
            self.envs.append(env)
            self.actions.append(a)

        add()
        self.footprint = self.replica_bytes(self.envs[0])
        r = max(2, -(-min_bytes // self.footprint))
        self.l2_note = "ring of %d replicas, %.0f MB total > 2x L2 (inputs larger than L2)"
        if r > max_replicas:
            r = max_replicas
            self.l2_note = "ring of %d replicas, %.0f MB total (SMALLER than L2: L2-resident run)"
        while len(self.envs) < r:
            add()
        self.l2_note = self.l2_note % (r, r * self.footprint / 1e6)
        torch.cuda.synchronize(device)

    @staticmethod
    def replica_bytes(env):
        p = env.pipeline
        ts = list(p.sim_tensors.values()) + list(p.state.values()) + [v for v in p.out.values() if v is not None]
        return sum(t.numel() * t.element_size() for t in ts)

    def __len__(self):
        return len(self.envs)

    def launches(self):
        return sum(e.pipeline.launches for e in self.envs)

    def capture(self, fn, count, start=0):
        """Capture `count` consecutive ring steps (replica start, start+1, ...) running fn(env, actions)."""
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(count):
                j = (start + i) % len(self.envs)
                fn(self.envs[j], self.actions[j])
        return g


WARM_MS = 25.0  # minimum untimed activity of the timed graphs before a timed region


def warm_plan(plan, steps_per_graph, warmup_steps, device):
    """Untimed replays of the SAME CUDAGraph objects that are timed afterwards: every graph of the plan runs at least
    once (a graph's first launch pays its one-time upload), and at least `warmup_steps` ring steps run in total."""
    done = 0
    for (g, _), spg in zip(plan, steps_per_graph):
        g.replay()
        done += spg
    while done < warmup_steps:
        plan[0][0].replay()
        done += steps_per_graph[0]
    # ... and at least WARM_MS of them: a GPU that has just been idle runs its first half millisecond of kernels 3-6 % slower
    # (every kernel, the empty one included), which is most of a 20-step timed region but nothing of a 4000-step one
    t0 = time.perf_counter()
    while (time.perf_counter() - t0) * 1e3 < WARM_MS:
        for _ in range(8):
            plan[0][0].replay()
        done += 8 * steps_per_graph[0]
        torch.cuda.synchronize(device)
    torch.cuda.synchronize(device)
    return done


def timed_replays(plan, device, dist_barrier=None):
    """plan = [(graph, replays)]; returns device milliseconds between two events on the current stream."""
    if dist_barrier is not None:
        dist_barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # ~100 us of device-side delay AHEAD of the start event: the host enqueues the event and the first graph while the GPU
    # is still busy, so the timed region holds the K steps back to back and not the host's launch latency of the first
    # graph (10 us: 2 % of a 20-step run, nothing of a 4000-step one)
    torch.cuda._sleep(200_000)
    e0.record()
    for g, reps in plan:
        for _ in range(reps):
            g.replay()
    e1.record()
    torch.cuda.synchronize(device)
    if dist_barrier is not None:
        dist_barrier()
    return e0.elapsed_time(e1)


def make_plan(ring, fn, steps, group):
    """Graphs that together run exactly `steps` ring steps: q replays of a `group`-step graph + one remainder graph.
    Returns (plan, steps_per_graph)."""
    plan, spg = [], []
    q, r = divmod(steps, group)
    if q:
        plan.append((ring.capture(fn, group), q))
        spg.append(group)
    if r:
        plan.append((ring.capture(fn, r, start=(q * group) % len(ring)), 1))
        spg.append(r)
    return plan, spg


def full_step(env, a):
    env.pipeline.step(a)


def post_only(env, a):
    env.pipeline.post_step()  # the post-physics launch exactly as the step runs it (epilogue fused into its last CTA)


def pre_only(env, a):
    env.pipeline.pre_physics(a)


def epi_only(env, a):
    env.pipeline.epilogue()


def floor_only(env, a):
    p = env.pipeline  # an empty kernel launched like dex_pre_physics: the per-launch floor in this stream of graphs
    p.lib.dex_debug_launch_floor(p.n, p._stream())


# ------------------------------------------------------------------------------------------------
# CPU baseline: the oracle (CPU restatement of the reference's torch components) on the host cores.
# The ONLY place bench.py touches oracle/.
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(task, n, steps, warmup, seed, budget_s=None):
    """Time `steps` control steps of the oracle at `n` envs (after `warmup`); if budget_s is given,
    run as many steps as fit (>= 3).  Returns (env_steps_per_s, ms_per_step, steps_done, threads)."""
    from oracle import synth as osynth
    from oracle.dexstep_oracle import OracleEnv
    from dexrobot_isaac_b200 import synthetic
    from dexrobot_isaac_b200.config import load_task_config
    from dexrobot_isaac_b200.params import BodyLayout

    lay = BodyLayout.default(with_box=task == "BlindGrasping")
    olay = osynth.BodyLayout(with_box=task == "BlindGrasping")
    env = OracleEnv(load_task_config(task, num_envs=n), n, olay, synthetic.dof_props())
    gen = torch.Generator().manual_seed(seed)
    ring = []
    for _ in range(4):  # pre-generated ring of states: generation is outside the timed region
        st = synthetic.random_state(n, lay, gen, "cpu")
        st["rigid_body_states"] = st.pop("rigid_body_state")
        ring.append((st, synthetic.random_actions(n, env.num_actions, gen), synthetic.random_uniforms(n, gen)))
    env.load_state(ring[0][0])
    env.reset(ring[0][2])

    def one(i):
        st, a, u = ring[i % len(ring)]
        env.pre_physics_step(a)
        env.load_state(st)  # stand-in for the physics step (a memcpy)
        env.post_physics_step(u)

    for i in range(max(1, warmup)):
        one(i)
    t0, done = time.perf_counter(), 0
    while True:
        one(done)
        done += 1
        el = time.perf_counter() - t0
        if budget_s is None:
            if done >= steps:
                break
        elif done >= 3 and (el >= budget_s or done >= steps):
            break
    return n * done / el, 1e3 * el / done, done, torch.get_num_threads()


def run_reference_arm(args, task, n_env, rank, world):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the reference is pure Python,
    needs Isaac Gym to instantiate and /root/reference does not travel) on ALL host cores, over the whole job's env
    count (n_env x world; a bounded sample of it when the run would take more than ~2 minutes).  Rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)  # torchrun exports OMP_NUM_THREADS=1: undo it for the CPU arm
    n_job = n_env * world
    total = args.steps + args.warmup
    _, ms_full, _, _ = cpu_reference_run(task, n_env, 2, 1, 1234)
    n_s = n_job
    while n_s > 64 and total * ms_full * (n_s / n_env) > 120e3:
        n_s //= 2
    value, ms, done, threads = cpu_reference_run(task, n_s, args.steps, args.warmup, 1234)
    sample = (f"{done} steps x {n_s} envs of {task} (job = {n_job} envs; oracle port of the reference's torch components, "
              f"{threads} threads; the port runs 1.05-2x faster than the unmodified reference components on the same cores, "
              f"BASELINE.md section 5)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "task": task, "envs_per_gpu": n_env, "envs_total": n_job,
                       "envs_per_step_sampled": n_s, "device": "cpu", "threads": threads},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def eager_gpu_reference(task, n, device, steps=12):
    """The reference's op chain as it runs in production - eager torch ON THE GPU (tasks/dexhand_base.py:893-942) -
    restated by the oracle port with every tensor on `device`.  Returns ms/step and CUDA kernel launches per step."""
    from oracle import synth as osynth
    from oracle.dexstep_oracle import OracleEnv
    from dexrobot_isaac_b200 import synthetic
    from dexrobot_isaac_b200.config import load_task_config
    from dexrobot_isaac_b200.params import BodyLayout

    lay = BodyLayout.default(with_box=task == "BlindGrasping")
    olay = osynth.BodyLayout(with_box=task == "BlindGrasping")
    gen = torch.Generator(device=device).manual_seed(7)
    cgen = torch.Generator().manual_seed(7)
    with torch.device(device):
        env = OracleEnv(load_task_config(task, num_envs=n), n, olay, synthetic.dof_props().to(device))
        ring = []
        for _ in range(3):
            st = synthetic.random_state(n, lay, gen, device)
            st["rigid_body_states"] = st.pop("rigid_body_state")
            ring.append((st, synthetic.random_actions(n, env.num_actions, gen, device),
                         synthetic.random_uniforms(n, cgen).to(device)))
        env.load_state(ring[0][0])
        env.reset(ring[0][2])

        def one(i):
            st, a, u = ring[i % len(ring)]
            env.pre_physics_step(a)
            env.load_state(st)
            env.post_physics_step(u)

        for i in range(3):
            one(i)
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            one(i)
        e1.record()
        torch.cuda.synchronize(device)
        wall_ms = 1e3 * (time.perf_counter() - t0) / steps
        dev_ms = e0.elapsed_time(e1) / steps
        launches = None
        try:
            from torch.profiler import ProfilerActivity, profile

            with profile(activities=[ProfilerActivity.CUDA]) as prof:
                one(0)
                torch.cuda.synchronize(device)
            launches = sum(1 for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA)
        except Exception:
            pass
    ms = max(dev_ms, wall_ms)
    return {"ms_per_step": ms, "device_ms_per_step": dev_ms, "value": n / (ms * 1e-3), "unit": UNIT,
            "cuda_launches_per_step": launches,
            "what": "oracle port of the reference's torch components, eager torch, all tensors on the same B200"}


# ------------------------------------------------------------------------------------------------
# end to end: host buffers in, host buffers out, copies inside the timed region
# ------------------------------------------------------------------------------------------------
def run_e2e(ring, steps, device, barrier, compact=False):
    envs = ring.envs[:2]
    # the pinned staging buffers go to the GPU's own NUMA node (first touch under a temporary CPU binding; multi-GPU runs
    # are bound for good in main()): where they land otherwise changes from run to run, and the copy rates with it
    before = os.sched_getaffinity(0)
    bind_to_gpu_numa_node(device)
    try:
        bufs = [e.alloc_host_buffers(compact=compact) for e in envs]
    finally:
        os.sched_setaffinity(0, before)
    for e, a, (hin, _) in zip(envs, ring.actions, bufs):
        e.fill_host_inputs(hin, a)
    main = torch.cuda.current_stream(device)
    h2d = d2h = 0

    def loop(k):  # step_host pipelines uploads, kernels and downloads over its own copy streams
        nonlocal h2d, d2h
        for i in range(k):
            j = i % len(envs)
            h2d, d2h = envs[j].step_host(*bufs[j])

    loop(20)
    torch.cuda.synchronize(device)
    if barrier is not None:
        barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(main)
    loop(steps)
    for e in envs:
        e.host_sync()  # the downloads run on the download stream
    e1.record(main)
    torch.cuda.synchronize(device)
    if barrier is not None:
        barrier()
    for e in envs:
        e.end_host_mode()
    return e0.elapsed_time(e1), h2d, d2h


def bind_to_gpu_numa_node(device):
    """Multi-GPU runs: pin this rank to the CPUs NVML reports as local to its GPU BEFORE any pinned host buffer is
    allocated, so that the end-to-end path's staging buffers live on the GPU's own NUMA node (first-touch) and the
    H2D / D2H copies of the ranks do not all cross the same socket link.  Best effort: returns a note for the JSON line."""
    try:
        import pynvml

        pynvml.nvmlInit()
        idx = torch.cuda._get_nvml_device_index(device) if hasattr(torch.cuda, "_get_nvml_device_index") else device.index
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "nvml reports no local cpus inside this process's affinity mask"
        os.sched_setaffinity(0, cpus)
        return f"rank bound to {len(cpus)} cpus local to gpu {idx}"
    except Exception as e:  # noqa: BLE001
        return f"not bound ({type(e).__name__})"


def kernel_avg_us(ring, fn, launches, group, device):
    plan, spg = make_plan(ring, fn, launches, group)
    warm_plan(plan, spg, 3, device)  # the graphs that are timed, not copies of them
    return 1e3 * timed_replays(plan, device) / launches


class HotRing:
    """One replica replayed back to back: its ~40 MB working set stays in the 126 MB L2 (the L2-hot line)."""

    def __init__(self, ring):
        self.envs, self.actions = ring.envs[:1], ring.actions[:1]

    def __len__(self):
        return 1

    capture = Ring.capture


def main():
    args = parse_args()
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    task, n_env, b_pre, b_post = WORKLOADS[args.workload]
    if args.envs:
        n_env = args.envs
    if args.impl == "reference":
        run_reference_arm(args, task, n_env, rank, world)
        return 0
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the step pipeline has no CPU fallback")
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    numa_note = bind_to_gpu_numa_node(device) if world > 1 else "staging buffers allocated under a temporary binding to the gpu's local cpus"
    barrier = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=device)
        barrier = lambda: dist.barrier(device_ids=[local])  # noqa: E731
    from dexrobot_isaac_b200 import _lib
    from dexrobot_isaac_b200 import stats as dex_stats

    lib = _lib.load()  # fails loudly when the CUDA library is missing
    ring = Ring(task, n_env, device, 1234 + rank, args.observer, obs_norm=bool(args.obs_norm), nb_hand=args.nb_hand)
    if args.tile_kernel:
        for e in ring.envs:
            e.pipeline.tile_kernel = True
    if args.single_pass:
        for e in ring.envs:
            e.pipeline.single_pass = True
    R = len(ring)
    group = R * max(1, -(-32 // R))
    # The graphs that are timed are captured once and ALSO run the warm-up (>= --warmup untimed steps, every graph of
    # the plan at least once): the first launch of a CUDA graph pays a one-time upload that must not be timed.
    v2_before = lib.dex_debug_v2_launches()
    plan, spg = make_plan(ring, full_step, args.steps, group)
    v2_in_plan = lib.dex_debug_v2_launches() - v2_before
    warm_plan(plan, spg, max(args.warmup, 3), device)
    for e in ring.envs:
        e.pipeline.epi["totals"].zero_()
        if e.pipeline.out["stat_totals"] is not None:
            e.pipeline.out["stat_totals"].zero_()
    side = torch.cuda.Stream(device)
    sampler = ClockSampler(torch.cuda._get_nvml_device_index(device) if hasattr(torch.cuda, "_get_nvml_device_index") else local)

    # ---- the timed region: exactly K steps, barrier + synchronize on both sides, device events ----
    if barrier is not None:
        barrier()
    torch.cuda.synchronize(device)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for g, reps in plan:
        for _ in range(reps):
            g.replay()
    e1.record()
    # the single statistics all-reduce of the logging interval, on a side stream; timed by itself
    side.wait_event(e1)
    with torch.cuda.stream(side):
        a0.record()
        payload = dex_stats.pack_pipelines([e.pipeline for e in ring.envs])
        dex_stats.all_reduce(payload)  # NCCL over NVLink when world > 1, no-op on one GPU
        a1.record()
    torch.cuda.synchronize(device)
    if barrier is not None:
        barrier()
    sampler.stop()
    ms = e0.elapsed_time(e1)
    allreduce_us = 1e3 * a0.elapsed_time(a1)
    if world > 1:
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    stats = payload.cpu().tolist()
    C = _lib.CONSTS
    value = n_env * world * args.steps / (ms * 1e-3)

    # ---- dominant-kernel timing for the roofline (post-physics launches only, same ring, graph replay) ----
    kl = max(R, min(args.steps, 2000))
    post_us = kernel_avg_us(ring, post_only, kl, group, device)
    pre_us = kernel_avg_us(ring, pre_only, kl, group, device)
    fused = ring.envs[0].pipeline.fuse_epilogue
    epi_us = None if fused else kernel_avg_us(ring, epi_only, kl, group, device)
    launches_per_step = 2 if fused else 3
    peak, peak_src = measured_peak_hbm()
    achieved = b_post * n_env / (post_us * 1e-6) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{task}_{n_env}")
    except Exception:
        pass
    post_kernel = ("post_physics_kernel", "post_physics_v2_kernel", "post_physics_v2s_kernel",
                   "post_physics_v2_kernel<TMA>", "post_physics_v2_kernel<pair>")[lib.dex_debug_last_post_kernel()]
    floor_us = kernel_avg_us(ring, floor_only, kl, group, device)
    roofline = {"bound": "hbm", "kernel": post_kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "profiles/traffic.json (static: dram__bytes_read + dram__bytes_write of one launch, ncu --set full)",
                "peak_source": peak_src, "algorithmic_bytes_per_env": b_post, "launch_us": post_us,
                "step_frac": (b_pre + b_post) * n_env * args.steps / (ms * 1e-3) / 1e9 / peak}

    # ---- L2-hot line (report both): ONE replica replayed back to back, working set inside the 126 MB L2 ----
    hot = HotRing(ring)
    hk = max(32, min(args.steps, 1000))
    hot_step = kernel_avg_us(hot, full_step, hk, 32, device)
    hot_post = kernel_avg_us(hot, post_only, hk, 32, device)
    l2_hot = {"step_us": hot_step, "post_us": hot_post, "value": n_env * world / (hot_step * 1e-6),
              "post_frac_of_hbm_peak": b_post * n_env / (hot_post * 1e-6) / 1e9 / peak,
              "note": "one replica replayed back to back (inputs stay in L2); the headline value is the L2-cold ring"}

    # ---- the other geometry of the post-physics pass, same run (default: persistent kernel; alternative: single pass,
    #      which only exists for BlindGrasping between 2 and 4 tiles per SM) ----
    alt_geometry = None
    if not args.tile_kernel and not args.single_pass:
        for e in ring.envs:
            e.pipeline.single_pass = True
        alt_post = kernel_avg_us(ring, post_only, kl, group, device)
        if lib.dex_debug_last_post_kernel() == 2:
            alt_step = kernel_avg_us(ring, full_step, kl, group, device)
            alt_geometry = {"kernel": "post_physics_v2s_kernel", "how": "DEX_FLAG_SINGLE_PASS / StepPipeline.single_pass",
                            "post_us": alt_post, "step_us": alt_step, "value": n_env * world / (alt_step * 1e-6),
                            "l2_hot_post_us": kernel_avg_us(hot, post_only, hk, 32, device),
                            "l2_hot_step_us": kernel_avg_us(hot, full_step, hk, 32, device),
                            "note": "slower than the default when every input comes from HBM, faster when the sim tensors are "
                                    "L2-resident (post-physics right behind the simulator)"}
        for e in ring.envs:
            e.pipeline.single_pass = False

    # ---- end to end through DexStepEnv.step_host (host buffers, copies inside the timed region) ----
    # its own step count (reported in e2e.steps): the three-stage pipeline of step_host needs ~2 steps to fill and to drain,
    # which a 20-step region would count as 10-25 % (tools/debug/host_step_overhead.py)
    ke = max(200, min(args.steps, 2000))
    ems, h2d, d2h = run_e2e(ring, ke, device, barrier, compact="packed")
    cms, ch2d, cd2h = run_e2e(ring, ke, device, barrier, compact=True)
    fms, fh2d, fd2h = run_e2e(ring, ke, device, barrier, compact=False)
    if world > 1:
        t = torch.tensor([ems, cms, fms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ems, cms, fms = (float(x) for x in t.tolist())
    e2e = {"value": n_env * world * ke / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": d2h, "steps": ke, "ms_per_step": ems / ke,
           "api": "DexStepEnv.step_host, packed host staging (only what the path reads: dof_state, the hand row, 10 tip / pad "
                  "poses as 7-float slots, the monitored contact rows, the box row, actions; one pinned block and one copy per direction; uploads, kernels and downloads on three streams)",
           "compact_rows_layout": {"value": n_env * world * ke / (cms * 1e-3), "h2d_bytes_per_step": ch2d,
                                   "d2h_bytes_per_step": cd2h, "ms_per_step": cms / ke,
                                   "api": "DexStepEnv.step_host, the monitored body rows shipped whole (13 floats each)"},
           "full_tensor_layout": {"value": n_env * world * ke / (fms * 1e-3), "h2d_bytes_per_step": fh2d,
                                  "d2h_bytes_per_step": fd2h, "ms_per_step": fms / ke,
                                  "api": "DexStepEnv.step_host, whole Isaac Gym tensors shipped"}}
    if numa_note:
        e2e["numa"] = numa_note
    clocks = sampler.summary()

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "task": task, "envs_per_gpu": n_env, "envs_total": n_env * world,
                       "obs_dim": ring.envs[0].num_observations, "num_bodies": ring.layout.num_bodies,
                       "observer": bool(args.observer), "obs_norm": bool(args.obs_norm), "l2": ring.l2_note, "graph_steps": group,
                       "timing": "CUDA events on the launching stream around exactly K steps of CUDA-graph replays (the graphs that ran "
                                 "the warm-up); a ~100 us device-side delay ahead of the start event keeps the host's launch latency "
                                 "of the first graph out of the region",
                       "early_sim_loads": os.environ.get("DEXSTEP_EARLY_SIM_LOADS", "1") != "0",
                       "parallelism": f"env-sharded x{world}, no step-path collective, 1 NCCL all-reduce of stats"},
            "roofline": roofline, "e2e": e2e, "gpu_launches": launches_per_step * args.steps, "clocks": clocks,
            "single_pass_geometry": alt_geometry, "kernels_us": {"pre_physics": pre_us, "post_physics": post_us, "epilogue": epi_us, "empty_kernel_launch_floor": floor_us,
                           "stats_allreduce": allreduce_us},
            "l2_hot": l2_hot,
            "episode_stats": {"steps": stats[C["DEX_NUM_COUNTERS"]], "resets": stats[C["DEX_CNT_RESET"]],
                              "success": stats[C["DEX_CNT_SUCCESS"]], "failure": stats[C["DEX_CNT_FAILURE"]],
                              "timeout": stats[C["DEX_CNT_TIMEOUT"]],
                              "episode_length_sum": stats[C["DEX_CNT_EPISODE_LENGTH"]]}}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        torch.set_num_threads(os.cpu_count() or 1)
        v, cms, done, threads = cpu_reference_run(task, n_env, 10**9, 2, 1234, budget_s=args.cpu_seconds)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "ms_per_step": cms,
                                "sample": f"{done} steps x {n_env} envs of {task} (same workload; oracle port of the "
                                          f"reference's torch components; os.cpu_count()={os.cpu_count()}; the port runs "
                                          f"1.05-2x faster than the unmodified reference, BASELINE.md section 5)"}
        torch.set_num_threads(1)
        v1, cms1, done1, _ = cpu_reference_run(task, n_env, 10**9, 1, 1234, budget_s=min(6.0, args.cpu_seconds))
        line["cpu_baseline_1thread"] = {"value": v1, "unit": UNIT, "cores": 1, "kind": "port", "ms_per_step": cms1,
                                        "sample": f"{done1} steps x {n_env} envs"}
        torch.set_num_threads(os.cpu_count() or 1)
        try:
            line["reference_gpu_eager"] = eager_gpu_reference(task, n_env, device)
        except Exception as ex:  # the line is informational: never let it take the bench down
            line["reference_gpu_eager"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}
    if args.sweep and rank == 0:
        sweep = []
        for n in (1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072, 262144):
            r = Ring(task, n, device, 99, args.observer)
            if args.tile_kernel:
                for e in r.envs:
                    e.pipeline.tile_kernel = True
            g = len(r) * max(1, -(-32 // len(r)))
            us = kernel_avg_us(r, post_only, max(len(r), 600), g, device)
            sus = kernel_avg_us(r, full_step, max(len(r), 600), g, device)
            sweep.append({"envs": n, "post_us": us, "post_frac": b_post * n / (us * 1e-6) / 1e9 / peak,
                          "step_us": sus, "env_steps_per_s": n / (sus * 1e-6), "replicas": len(r)})
            del r
            torch.cuda.empty_cache()
        line["sweep"] = sweep
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
