#!/usr/bin/env python
"""bench.py -- agent-steps/s of the batched CollisionAvoidanceEnv step on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload full|c2|c3|c4|c5|rvo] [--impl reference]

One "step" = one pass of the hot path (casim_step, preceded by the policy kernel of the workload) over one batch of E envs
x A agents of synthetic scenarios (casim/scenarios.py).  DEFAULT WORKLOAD = `full`: the configuration
BASELINE.json's north_star quotes its target on -- 4096 envs x 10 agents per GPU with OtherAgentsStates + LaserScan +
OccupancyGrid on the crossing map.  Prints ONE JSON line (rank 0).

Timing: device-side CUDA events on the launching stream around `reps` replays of ONE CUDA graph of exactly K steps
(reps chosen so that the timed region is >= 50 ms; ms_per_step = time / (K * reps)), after W warm-up steps, with a
barrier + synchronize on both sides, max over ranks.  The K steps walk a RING of independent env batches whose combined
footprint exceeds the 126 MB L2 ("inputs larger than L2").  The design's one collective -- the all-reduce of the [8]
per-step episode statistics -- is issued once per step INSIDE the graph on a side stream (NCCL, N > 1), overlapping the
next step's kernels.  `e2e` is the same metric through the host-buffer call (pinned host actions in, results and
observations out) per step.  `cpu_baseline` / `--impl reference` time the CPU oracle port (oracle/casim_oracle.c, plain
C fp64 restatement of the reference's Python step, POSIX threads over environments) on the host cores: the reference
itself is Python that needs /root/reference, which does not exist on the GPU box; its own step rate, measured in the
build container by oracle/time_reference_python.py, is quoted from profiles/reference_python_rates.json.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "gym-exploration-2d_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

L2_BYTES = 126 * 1024 * 1024

# algorithmic bytes per agent-step (DESIGN.md "Algorithmic bytes")
BYTES_CORE = 177          # 96 B state read + 76 B mutable state written + 4 B reward + 1 B done
BYTES_OAS = 384           # 9 x 10 fp32 rows + 6 scalar observation words
BYTES_LASER = 320         # 16 fp32 ranges written + 256 one-byte cell samples
BYTES_WALL = 81           # disc of cells around the agent (r = 0.5 m)
BYTES_GRID = 11169        # 3600 B written + <= 87^2 B rotated window read
# bytes that HAVE to cross HBM per agent-step: what is written plus the state that is read; a shared map
# lives in L2 / shared memory and costs nothing, a per-env map adds its bit planes once per environment
MUST_CORE = 177
MUST_OAS = 384
MUST_LASER = 64
MUST_GRID = 3600
MAP_PLANE_BYTES = 4 * 300 * 40     # four footprint bit planes of one map (per-env-map runs only)

WORKLOADS = {
    "c2": dict(E=4096, A=10, sensors=(), desc="4096 envs x 10 agents, NonCooperative/Static mix, collision + reward step only"),
    "c3": dict(E=4096, A=10, sensors=("other_agents_states", "laserscan"),
               desc="4096 envs x 10 agents, OtherAgentsStates + LaserScan ray-casting against the crossing map"),
    "full": dict(E=4096, A=10, sensors=("other_agents_states", "laserscan", "local_grid"),
                 desc="4096 envs x 10 agents, OtherAgentsStates + LaserScan + OccupancyGrid, crossing map"),
    "c4": dict(E=1024, A=4, sensors=("other_agents_states",),
               desc="1024 envs x 4 agents (2 ig_greedy + 2 Static targets), exploration: target-map odds update + "
                    "information-gain scoring of 9 candidate poses, crossing map"),
    "c5": dict(E=1024, A=20, sensors=("other_agents_states",),
               desc="8192/8 = 1024 envs per GPU x 20 agents (CADRL formations), batched GA3C-CADRL LSTM policy forward, "
                    "random-init weights"),
    "rvo": dict(E=4096, A=10, sensors=(),
                desc="4096 envs x 10 agents, all RVOPolicy (ORCA, reciprocal), collision + reward step (SURVEY 8(f) rank 1)"),
}
# 2*[T*71*256 + 68*256 + 256*256 + 256*256 + 256*11] with T = LSTM steps actually run: Config.MAX_NUM_OTHER_AGENTS_OBSERVED = 9
# (SURVEY.md 8(d) counts the network's 10 slots: 666 112; the 10th slot is never filled)
GA3C_FLOP_PER_AGENT = 2 * (9 * 71 * 256 + 68 * 256 + 256 * 256 + 256 * 256 + 256 * 11)


def algorithmic_bytes(wl):
    b = BYTES_CORE
    s = WORKLOADS[wl]["sensors"]
    if "other_agents_states" in s:
        b += BYTES_OAS
    if "laserscan" in s:
        b += BYTES_LASER + BYTES_WALL
    if "local_grid" in s:
        b += BYTES_GRID
    return b


def must_move_bytes(wl, per_env_maps=False, A=10):
    b = MUST_CORE
    s = WORKLOADS[wl]["sensors"]
    if "other_agents_states" in s:
        b += MUST_OAS
    if "laserscan" in s:
        b += MUST_LASER
    if "local_grid" in s:
        b += MUST_GRID
    if per_env_maps and ("laserscan" in s or "local_grid" in s):
        b += MAP_PLANE_BYTES / A
    return b


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["bf16_tflops"]), "measured (MEASURED_PEAKS.json bf16_tflops, burst: kernel timed alone)"
    except Exception:
        return 2250.0, "fallback (nominal dense bf16)"


def roofline_obj(wl, kern, dominant, step_s, step_bytes, step_must, explore_counts):
    """Roofline of the DOMINANT kernel of the step (algorithmic bytes or flops per launch / its measured launch
    duration) plus the per-kernel breakdown and the whole-step figure.  Two byte counts are reported for HBM-bound
    kernels: `frac` on SURVEY.md 8(d)'s algorithmic bytes (which count the rotated map window of the occupancy grid and
    the laser's cell samples as if they came from HBM) and `frac_must_move` on the bytes that have to cross HBM."""
    peak, peak_src = peaks()
    t, units, how, must = kern[dominant]
    if dominant == "k_ga3c_fwd(fix-up)":
        # exact mode: the network's algorithmic flops are produced by the fp16-operand tensor-core forward PLUS the fp32
        # re-evaluation of the agents it cannot call; the roofline line is the pair, on the flops of ONE forward per agent
        t_tc, units, _, _ = kern["k_ga3c_tc"]
        t, how, dominant = t + t_tc, "k_ga3c_tc timed alone + (exact mode minus the bare tensor-core forward)", "k_ga3c_tc + k_ga3c_fwd(fix-up)"
    tensor = dominant.startswith(("k_ga3c_tc", "k_ga3c_fwd"))
    if tensor:
        tpeak, tsrc = tensor_peak()
        ach = units / t / 1e12
        out = {"bound": "tensor", "kernel": dominant, "achieved": ach, "peak": tpeak, "unit": "TFLOP/s", "frac": ach / tpeak,
               "peak_source": tsrc, "algorithmic_flops_per_launch": units}
    else:
        ach = units / t / 1e9
        out = {"bound": "hbm", "kernel": dominant, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
               "peak_source": peak_src, "algorithmic_bytes_per_launch": units,
               "must_move_bytes_per_launch": must, "achieved_must_move": must / t / 1e9, "frac_must_move": must / t / 1e9 / peak}
    out["duration_us"] = t * 1e6
    out["how"] = how
    out["traffic"] = traffic_for(wl, dominant)
    out["kernels"] = {k: {"duration_us": v[0] * 1e6, "share_of_step": v[0] / step_s, "how": v[2],
                          "frac": (v[1] / v[0] / 1e9 / peak) if not k.startswith("k_ga3c") else None,
                          "frac_must_move": (v[3] / v[0] / 1e9 / peak) if not k.startswith("k_ga3c") else None,
                          "tensor_frac": (v[1] / v[0] / 1e12 / tensor_peak()[0]) if k in ("k_ga3c_tc", "k_ga3c_fwd") else None}
                      for k, v in kern.items()}
    out["whole_step"] = {"duration_us": step_s * 1e6, "algorithmic_bytes": step_bytes,
                         "achieved_GBps": step_bytes / step_s / 1e9, "frac_of_hbm_peak": step_bytes / step_s / 1e9 / peak,
                         "must_move_bytes": step_must, "frac_must_move": step_must / step_s / 1e9 / peak}
    if explore_counts is not None:
        out["explore_counters_one_launch"] = dict(zip(("cells_tested", "edf_samples", "visible_cells", "clamped"), explore_counts))
    return out


def traffic_for(wl, kernel=None):
    """dram bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f).get(wl)
        if not isinstance(d, dict):
            return d
        key = (kernel or "").split("(")[0].split(" ")[0]   # "k_grid_tma(grid+laser)" -> "k_grid_tma"
        return d.get(key, d.get({"k_rvo": "k_rvo4"}.get(key, key)))
    except Exception:
        return None


def reference_python_rates(wl):
    """The reference's own Python step rate for this workload, measured in the build container (one core) by
    oracle/time_reference_python.py -- it cannot run on the GPU box (/root/reference is absent there)."""
    try:
        with open(os.path.join(ROOT, "profiles", "reference_python_rates.json")) as f:
            d = json.load(f)
        w = dict(d["workloads"].get(wl, {}))
        w["measured"] = "build container, 1 core (%s), %s" % (d.get("host", "?"), d.get("how", ""))
        return w
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
_C2_CASES = None


def c2_cases():
    """The reference's 10-agent scenario suite (tests/golden/c2_10_agents_500_cases.npz, derived from
    envs/test_cases/10_agents_500_cases.p by oracle/gen_bench_scenarios.py)."""
    global _C2_CASES
    if _C2_CASES is None:
        _C2_CASES = np.load(os.path.join(ROOT, "tests", "golden", "c2_10_agents_500_cases.npz"))["cases"]
    return _C2_CASES


def scenario(wl, E, A, seed):
    from casim import layout as L
    from casim import scenarios as S
    sensors = WORKLOADS[wl]["sensors"]
    mask = 0
    if "other_agents_states" in sensors:
        mask |= L.SENSOR_OTHER_AGENTS
    if "laserscan" in sensors:
        mask |= L.SENSOR_LASERSCAN
    if "local_grid" in sensors:
        mask |= L.SENSOR_OCCUPANCY_GRID
    if wl == "c2":
        init, types, n = (S.from_cases(c2_cases(), E, A, seed=seed) if A == 10 else S.open_space(E, A, seed=seed, sensors=0))
        return init, types, n, None
    if wl == "rvo":
        if A == 10:
            init, types, n = S.from_cases(c2_cases(), E, A, seed=seed, static_every=0, policy=L.POLICY_RVO)
        else:
            init, types, n = S.open_space(E, A, seed=seed, sensors=0, static_every=0, policy=L.POLICY_RVO)
        return init, types, n, None
    if wl == "c4":
        return S.exploration(E, seed=seed)
    if wl == "c5":
        init, types, n = S.formations(E, A, seed=seed)
        return init, types, n, None
    return S.crossing(E, A, seed=seed, sensors=mask)


def workload_config(wl, E, A, world, per_env_maps=False):
    """The `config` object of the JSON line -- identical for the GPU arm and the reference arm (it names the workload,
    not the implementation)."""
    W = WORKLOADS[wl]
    return {"workload": wl + ": " + W["desc"], "envs_per_gpu": E, "agents_per_env": A, "sensors": list(W["sensors"]),
            "scenario": SCENARIO_DESC[wl], "map": ("per-env maps" if per_env_maps else "one map shared by all envs") if wl in ("c3", "full", "c4") else "none",
            "auto_reset": True, "n_ranks": world,
            "l2": "inputs larger than L2: ring of independent env batches, > 1.5 x 126 MB resident per GPU",
            "launch": "one CUDA graph of K steps, replayed `reps` times (timed region >= 50 ms)",
            "parallelism": "envs sharded over the ranks, no data-path collective; per-step all-reduce of the [8] episode "
                           "statistics on a side stream inside the timed region (NCCL when n_ranks > 1)",
            "state": "fp32 SoA state, fp64 predicates (dtype = arithmetic of the decision path)"}


SCENARIO_DESC = {
    "c2": "env e = case (e mod 500) of the reference's 10_agents_500_cases suite + N(0, 0.05^2) jitter; agents i%5==4 Static",
    "rvo": "same cases as c2, every agent RVOPolicy",
    "c3": "starts in the free cross of the crossing map, arm uniform, along U(3,12), across U(-1.4,1.4), goal = -start, starts >= 1.2 m apart",
    "full": "starts in the free cross of the crossing map, arm uniform, along U(3,12), across U(-1.4,1.4), goal = -start, starts >= 1.2 m apart",
    "c4": "2 ig_greedy explorers on the cross + 2 static targets (IG_agent_crossing layout)",
    "c5": "starts on a radius-10 circle, goals along the strokes of the formation() letter 'CADRL'[e mod 5]",
}


def ga3c_trained_weights():
    """The reference's shipped GA3C-CADRL checkpoint (GA3C_CADRL/checkpoints/IROS18/network_01900000, the default of
    GA3CCADRLPolicy.initialize_network), as read from the TF1 bundle by casim.ga3c_checkpoint and committed as a test
    fixture; None when the fixture is absent."""
    path = os.path.join(ROOT, "tests", "golden", "ga3c_iros18.npz")
    return np.load(path)["weights"] if os.path.exists(path) else None


def ga3c_weights(seed=0):
    """Random-init weights of the GA3C-CADRL architecture: Glorot-uniform kernels, zero biases (TF1 defaults)."""
    rng = np.random.default_rng(seed)
    shapes = [(71, 256), (256,), (68, 256), (256,), (256, 256), (256,), (256, 256), (256,), (256, 11), (11,), (256, 1), (1,)]
    parts = []
    for sh in shapes:
        if len(sh) == 2:
            lim = np.sqrt(6.0 / (sh[0] + sh[1]))
            parts.append(rng.uniform(-lim, lim, sh).astype(np.float32).ravel())
        else:
            parts.append(np.zeros(sh, np.float32))
    return np.concatenate(parts)


def oracle_run(wl, E, A, steps, warmup, threads, seed=1234, per_env_maps=False):
    """CPU oracle port on the same synthetic workload; returns (agent-steps/s, seconds per step)."""
    from oracle import oracle
    cfg = oracle.default_config()        # byte-identical to casim_default_config (tests/test_abi_cpu.py); the CPU arms
    cfg.auto_reset = 1                   # never load libcasim.so
    init, types, n, rects = scenario(wl, E, A, seed)
    maps = None
    if rects is not None:
        maps = oracle.rasterize(cfg, rects.astype(np.float64))
        if per_env_maps:
            from casim.scenarios import per_env_crossing_rects
            maps = np.stack([oracle.rasterize(cfg, r.astype(np.float64)) for r in per_env_crossing_rects(E, seed=77)])
    # the oracle only runs sensors whose output buffer exists; give it exactly the workload's
    sensors = WORKLOADS[wl]["sensors"]
    state, obs = oracle.reset(cfg, init.astype(np.float64), types, n_agents=n, maps=maps)
    if "other_agents_states" not in sensors:
        obs.other_agents_states = None
    if "laserscan" not in sensors:
        obs.laserscan = None
    if "local_grid" not in sensors:
        obs.local_grid = None
    init64 = init.astype(np.float64)
    go = np.zeros((E,), np.uint8)
    belief = np.ones((E, A, cfg.ig_map_dim, cfg.ig_map_dim)) if wl == "c4" else None
    w = None
    if wl == "c5":
        w = ga3c_trained_weights()
        w = ga3c_weights() if w is None else w
    if wl == "c5":
        cfg.near_goal_threshold = 0.2

    def one_step():
        pol = None
        if wl == "c4":
            pol = oracle.explore_step(cfg, state, obs.other_agents_states, belief, maps, n_agents=n, init=init64,
                                      prev_game_over=go, n_threads=threads)["policy_actions"]
        elif wl == "c5":
            pol = oracle.ga3c_forward(cfg, state, w, n_agents=n, n_threads=threads)["policy_actions"]
        elif wl == "rvo":
            pol = oracle.rvo_actions(cfg, state, n_agents=n, n_threads=threads)
        oracle.step(cfg, state, policy_actions=pol, n_agents=n, maps=maps, init=init64, types0=types, prev_game_over=go,
                    obs=obs, n_threads=threads)
        return obs.game_over.copy()

    for _ in range(warmup):
        go = one_step()
    t0 = time.perf_counter()
    for _ in range(steps):
        go = one_step()
    dt = time.perf_counter() - t0
    return E * A * steps / dt, dt / max(steps, 1)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled in the background while the GPU is under load."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.perf_counter(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], [], set()
        for t, line in self.samples:
            if t < t0 or t > t1:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
REF_SAMPLE_ENVS = {"c2": 4096, "c3": 512, "full": 64, "c4": 128, "c5": 256, "rvo": 2048}


def cpu_baseline_obj(wl, v, threads, sample):
    out = {"value": v, "unit": "agent-steps/s", "cores": threads, "kind": "port", "sample": sample,
           "what": "oracle/casim_oracle.c: plain-C fp64 restatement of the reference's Python step (pinned to it by "
                   "tests/golden), POSIX threads over environments -- a STRONGER baseline than the reference's own Python"}
    rp = reference_python_rates(wl)
    if rp is not None:
        out["reference_python"] = rp
    return out


def run_reference(args, rank, world):
    """--impl reference: the CPU implementation of the path (oracle port, all host threads) on a bounded sample."""
    if rank != 0:
        return
    wl = args.workload
    W = WORKLOADS[wl]
    threads = min(os.cpu_count() or 1, 64)
    E, A = (args.envs or W["E"]), W["A"]
    # bounded sample: the full per-GPU batch for the cheap workloads, fewer envs when sensors make a step slow
    Es = min(REF_SAMPLE_ENVS[wl], E)
    steps = max(1, min(args.steps, 200 if wl == "c2" else 40))
    warm = max(1, min(args.warmup, 5))
    v, sps = oracle_run(wl, Es, A, steps, warm, threads, per_env_maps=bool(args.per_env_maps and wl in ('c3', 'full')))
    sample = "%d steps of %d envs x %d agents per step (oracle/casim_oracle.c, %d POSIX threads over envs)" % (steps, Es, A, threads)
    line = {
        "impl": "reference", "metric": "agent-steps/s", "value": v, "unit": "agent-steps/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(wl, E, A, world, args.per_env_maps),
        "cpu_baseline": cpu_baseline_obj(wl, v, threads, sample),
        "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


_JSON_FD = None


def guard_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to fd 1 when the box
    sets NCCL_DEBUG=VERSION), so fd 1 is pointed at stderr for the whole run and the line goes to the saved descriptor."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="full", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="casim", choices=["casim", "reference"])
    ap.add_argument("--envs", type=int, default=None, help="override envs per GPU (parity/sweep use; default = the named config)")
    ap.add_argument("--per-env-maps", action="store_true", help="c3/full: one map per environment (n_maps == E) instead of a shared map")
    ap.add_argument("--no-graph", action="store_true", help="launch every step from Python instead of one CUDA graph")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-collective", action="store_true", help="diagnostics: leave the per-step statistics all-reduce out")
    ap.add_argument("--min-timed-ms", type=float, default=50.0, help="replay the K-step graph until the timed region is this long")
    ap.add_argument("--e2e-steps", type=int, default=200)
    ap.add_argument("--ga3c", default="tc-exact", choices=["tc-exact", "tc", "fp32"],
                    help="c5 forward: tensor cores (fp16 operands, fp32 accumulate) + fp32 re-evaluation of the agents they cannot call (default: actions equal "
                         "the fp32 forward's), bare tensor cores, or the fp32 SIMT kernel")
    ap.add_argument("--ga3c-weights", default="iros18", choices=["iros18", "random"],
                    help="c5: the reference's shipped IROS18 checkpoint (default) or Glorot random-init weights")
    ap.add_argument("--no-stats", action="store_true", help="diagnostics: do not compute the per-step episode statistics")
    ap.add_argument("--steps-per-launch", type=int, default=1, help="c2 / c3 / full: casim_step_n, this many env steps per "
                    "kernel launch with the state in registers (in-kernel policies); K must be a multiple of it")
    ap.add_argument("--ring", type=int, default=None, help="override the number of env batches in the ring (diagnostics)")
    args = ap.parse_args()
    guard_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from casim import lib
    from casim.sim import BatchedSim
    from casim.config import NUM_STATS

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    wl = args.workload
    W = WORKLOADS[wl]
    E, A = (args.envs or W["E"]), W["A"]
    K, Wm = args.steps, max(args.warmup, 3)
    cfg = lib.default_config()
    cfg.auto_reset = 1
    per_env_maps = bool(args.per_env_maps and wl in ("c3", "full"))

    # ---- ring of independent env batches, footprint > L2 ----
    per_agent_resident = 96 + 5 + (16 if W["sensors"] else 0)
    if "other_agents_states" in W["sensors"]:
        per_agent_resident += 360
    if "laserscan" in W["sensors"]:
        per_agent_resident += 64 + 32
    if "local_grid" in W["sensors"]:
        per_agent_resident += 3600 + 32
    if wl == "c4":
        per_agent_resident += 60 * 60 * 4 + 8          # fp32 belief + action
    if wl == "c5":
        per_agent_resident += 76 * 4 + 8               # network input scratch + action
        cfg.near_goal_threshold = 0.2                   # experiments/src/run_cadrl_formations.py:21
    batch_bytes = E * A * per_agent_resident
    R = int(min(max(2, -(-int(1.5 * L2_BYTES) // batch_bytes)), 96))
    if args.ring:
        R = args.ring
    stats_rows = torch.zeros((max(K, Wm, 1), NUM_STATS), dtype=torch.float32, device=dev)
    sims = []
    for r in range(R):
        sim = BatchedSim(E, A, cfg=cfg, sensors=W["sensors"], want_debug=False, want_misc=bool(W["sensors"]), want_stats=not args.no_stats)
        init, types, n, rects = scenario(wl, E, A, seed=1234 + 1000 * rank + r)
        if rects is not None:
            if per_env_maps:
                from casim.scenarios import per_env_crossing_rects
                sim.set_obstacles(per_env_crossing_rects(E, seed=77 + r))
            else:
                sim.set_obstacles(rects)
        if wl == "c4":
            sim.enable_exploration(want_debug=False)
        if wl == "c5":
            wts = ga3c_trained_weights() if args.ga3c_weights == "iros18" else None
            sim.ga3c_set_weights(wts if wts is not None else ga3c_weights(), want_debug=False)
        sim.reset(init, types, n)
        sims.append(sim)
    torch.cuda.synchronize()
    use_tc = args.ga3c != "fp32"
    ga3c_exact = args.ga3c == "tc-exact"

    spl = max(1, int(args.steps_per_launch))
    if spl > 1:
        assert wl in ("c2", "c3", "full") and K % spl == 0 and Wm % spl == 0, "--steps-per-launch: c2/c3/full, K and W multiples of it"
    stream = torch.cuda.Stream(device=dev)
    from casim.dist import StatsReducer
    reducer = StatsReducer(device=dev)
    side = reducer.stream
    collective = (not args.no_collective) and (not args.no_stats)
    stats_events = []
    if collective:                    # the statistics are final right after k_core: the library records this event there
        for sm in sims:
            ev = torch.cuda.Event()
            ev.record(stream)
            sm.set_stats_event(ev)
            stats_events.append(ev)

    def policy_step(sim):
        if wl == "c4":
            return sim.explore_step()
        if wl == "c5":
            return sim.ga3c_forward(use_tensor_cores=use_tc, exact=ga3c_exact)
        if wl == "rvo":
            return sim.rvo_step()
        return None

    def run_steps(k0, k, reduce_stats=True):
        """Steps k0 .. k0+k-1 on `stream` (current).  Step i writes its episode statistics into row i of `stats_rows`
        (one row per step of the graph, so nothing has to be cloned) and the row is all-reduced across the ranks on
        the side stream while the next step's kernels run."""
        if spl > 1:                       # casim_step_n: spl steps per launch, each launch on the next ring batch
            for j, i in enumerate(range(k0, k0 + k, spl)):
                sim = sims[j % R]
                ring = sim.step_n(spl)
                if collective and reduce_stats:
                    reducer.reduce_row(ring["stats"], ready_event=stats_events[j % R])
            if collective and reduce_stats:
                reducer.wait()
            return
        for i in range(k0, k0 + k):
            sim = sims[i % R]
            row = stats_rows[i % stats_rows.shape[0]]
            if not args.no_stats:
                sim._obs.stats = row.data_ptr()
            sim.step(policy_actions=policy_step(sim))
            if collective and reduce_stats:            # NCCL all-reduce of this step's row, overlapping its sensor kernels
                reducer.reduce_row(row, ready_event=stats_events[i % R])
        if collective and reduce_stats:
            reducer.wait()

    with torch.cuda.stream(stream):
        run_steps(0, Wm)                                  # warm-up (also loads the kernels, creates the NCCL communicator)
        n0 = sims[0].launch_count()                       # kernels per step: counted by the library, not assumed
        run_steps(0, spl)
        launches_per_step = (sims[0].launch_count() - n0) / spl
    stream.synchronize()
    side.synchronize()

    graph = None
    if not args.no_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            with torch.cuda.graph(graph, stream=stream):
                run_steps(0, K)
        stream.synchronize()

    def one_pass():
        if graph is not None:
            graph.replay()
        else:
            run_steps(0, K)

    # untimed passes: graph upload, every ring batch touched once more, and the duration that sizes `reps`
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        one_pass()
        ev0.record(stream)
        one_pass()
        ev1.record(stream)
    stream.synchronize()
    pass_ms = max(ev0.elapsed_time(ev1), 1e-3)
    reps = max(1, int(-(-args.min_timed_ms // pass_ms)))
    if world > 1:                                        # every rank must replay the same number of collectives
        t = torch.tensor([reps], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        reps = int(t.item())

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.15)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- timed region: reps x exactly K steps ----
    barrier()
    t_load0 = time.perf_counter()
    with torch.cuda.stream(stream):
        ev0.record(stream)
        for _ in range(reps):
            one_pass()
        ev1.record(stream)
    barrier()
    ms = ev0.elapsed_time(ev1)
    stats_last = stats_rows[:K].clone() if spl == 1 else torch.cat([sm._ring["stats"] for sm in sims[:max(1, min(R, K // spl))]])
    # keep the same workload running ~1.2 s so that nvidia-smi (50 ms period) sees the clocks under this load
    t_end = time.perf_counter() + 1.2
    with torch.cuda.stream(stream):
        while time.perf_counter() < t_end:
            one_pass()
            stream.synchronize()
    t_load1 = time.perf_counter()
    if sampler:
        time.sleep(0.1)
        sampler.stop()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    total_agent_steps = float(E * A * K * reps * world)
    value = total_agent_steps / (ms * 1e-3)

    # ---- e2e: the same metric through the host-buffer calls of the C ABI, pinned host buffers, copies inside the timed
    # region.  Workloads whose policy runs in k_core and that carry sensor planes (c3, full) go through the PIPELINED
    # call: every result and observation reaches host memory every step, the occupancy grid as numpy.packbits bytes,
    # D2H of step t overlapping the kernels of step t+1 (the call returns step t-1: one step of latency).  The others
    # use the blocking call (their policy kernel reads the device observation of the same step). ----
    Ke = max(1, min(args.e2e_steps, 2000))
    fields = ["rewards", "done", "game_over"] + ([] if args.no_stats else ["stats"]) + list(W["sensors"]) + (["obs_misc"] if W["sensors"] else [])
    pipelined = bool(W["sensors"]) and wl in ("c3", "full")
    if pipelined:
        pipes = [sm.host_pipeline(fields, packed_grid=True, ext_actions=True, stream=stream) for sm in sims[:1]]
        pipe = pipes[0]
        h2d, d2h = pipe.h2d_bytes, pipe.d2h_bytes
        with torch.cuda.stream(stream):
            for i in range(3):
                pipe.step()
            pipe.flush()
            barrier()
            t0 = time.perf_counter()
            for i in range(Ke):
                pipe.step()
            pipe.flush()
            e2e_s = time.perf_counter() - t0
            barrier()
        e2e_call = ("casim_step_host_pipelined via BatchedSim.host_pipeline: pinned host buffers, actions read from mapped host "
                    "memory, scalars stored zero-copy, OtherAgentsStates / LaserScan planes + np.packbits occupancy grid copied "
                    "D2H on a copy stream overlapping the next step; results delivered with one step of latency")
    else:
        host = {"rewards": torch.zeros((E, A), dtype=torch.float32).pin_memory(),
                "done": torch.zeros((E, A), dtype=torch.uint8).pin_memory(),
                "game_over": torch.zeros((E,), dtype=torch.uint8).pin_memory(),
                "stats": torch.zeros((NUM_STATS,), dtype=torch.float32).pin_memory()}
        if args.no_stats:
            del host["stats"]
        for name in W["sensors"]:
            host[name] = torch.zeros(sims[0].buf[name].shape, dtype=sims[0].buf[name].dtype).pin_memory()
        if W["sensors"]:
            host["obs_misc"] = torch.zeros((E, A, 4), dtype=torch.float32).pin_memory()
        act_host = torch.zeros((E, A, 2), dtype=torch.float32).pin_memory()
        act_dev = torch.zeros((E, A, 2), dtype=torch.float32, device=dev)
        h2d = act_host.numel() * 4
        d2h = sum(t.numel() * t.element_size() for t in host.values())
        host = BatchedSim.HostView(host)
        with torch.cuda.stream(stream):
            # pre-bound host steps (one per ring batch): all ctypes arguments resolved once, like an RL loop would
            pa = [None] * R
            if wl in ("c4", "c5", "rvo"):
                for s_ in sims:
                    policy_step(s_)                                  # allocates buf["policy_actions"]
                pa = [s_.buf["policy_actions"] for s_ in sims]
            host_step = [sims[r].bind_host_step(host, ext_actions_host=act_host, ext_actions_dev=act_dev, policy_actions=pa[r],
                                                stream=stream) for r in range(R)]
            for i in range(3):
                policy_step(sims[i % R])
                host_step[i % R]()
            barrier()
            ev0.record(stream)
            t0 = time.perf_counter()
            for i in range(Ke):
                policy_step(sims[i % R])
                host_step[i % R]()
            ev1.record(stream)
            barrier()
            e2e_s = max(ev0.elapsed_time(ev1) * 1e-3, time.perf_counter() - t0)
        e2e_call = "casim_step_host via BatchedSim.bind_host_step (blocking, pinned host buffers; scalars zero-copy)"
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = float(E * A * Ke * world) / e2e_s

    # ---- episode statistics of the timed region's last pass (rows already all-reduced inside the graph when N > 1) ----
    st = stats_last.double().cpu().numpy()
    if world > 1 and not collective and not args.no_stats:
        t = stats_last.clone()
        dist.all_reduce(t)
        st = t.double().cpu().numpy()
    n_ag = max(float(st[:, 6].mean()), 1.0)
    episode = {"per_step_mean": {"new_collisions": float(st[:, 0].mean()), "new_at_goal": float(st[:, 1].mean()),
                                 "new_timeouts": float(st[:, 2].mean()), "game_over_envs": float(st[:, 3].mean()),
                                 "reward_sum": float(st[:, 4].mean()), "active_agents": float(st[:, 5].mean()),
                                 "agents": float(st[:, 6].mean())},
               "active_agent_fraction": float(st[:, 5].mean()) / n_ag,
               "new_collision_fraction_per_step": float(st[:, 0].mean()) / n_ag,
               "env_resets_per_step": float(st[:, 3].mean()),
               "all_ranks": world > 1}

    # ---- per-kernel durations, measured live with CUDA events on the launching stream: each kernel class is
    # replayed alone (one CUDA graph of `reps` launches walking the same ring of batches, inputs > L2) ----
    from casim import layout as LY

    def time_alone(fn, reps=200):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(stream):
            for i in range(3):
                fn(sims[i % R])
            stream.synchronize()
            with torch.cuda.graph(g, stream=stream):
                for i in range(reps):
                    fn(sims[i % R])
            g.replay()
            stream.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            g.replay()
            b.record(stream)
            stream.synchronize()
        return a.elapsed_time(b) * 1e-3 / reps

    step_s = ms * 1e-3 / (K * reps)
    kern = {}          # name -> (seconds per launch, algorithmic units per launch, how)
    NA = E * A
    t_sens = 0.0
    fused = "laserscan" in W["sensors"] and "local_grid" in W["sensors"] and not per_env_maps
    if fused:
        # shared map: ONE persistent kernel (k_grid_tma) stages the map in shared memory and produces both observations
        t = time_alone(lambda sm: sm.run_sensor(LY.SENSOR_LASERSCAN | LY.SENSOR_OCCUPANCY_GRID))
        kern["k_grid_tma(grid+laser)"] = (t, (BYTES_GRID + BYTES_LASER) * NA, "timed alone", (MUST_GRID + MUST_LASER) * NA)
        t_sens += t
    else:
        if "laserscan" in W["sensors"]:
            t = time_alone(lambda sm: sm.run_sensor(LY.SENSOR_LASERSCAN))
            kern["k_laser"] = (t, BYTES_LASER * NA, "timed alone", MUST_LASER * NA + (MAP_PLANE_BYTES * E if per_env_maps else 0))
            t_sens += t
        if "local_grid" in W["sensors"]:
            t = time_alone(lambda sm: sm.run_sensor(LY.SENSOR_OCCUPANCY_GRID))
            kern["k_grid"] = (t, BYTES_GRID * NA, "timed alone", MUST_GRID * NA + (MAP_PLANE_BYTES * E if per_env_maps else 0))
            t_sens += t
    explore_counts = None
    if wl == "c4":
        t = time_alone(lambda sm: sm.explore_step())
        cnt = torch.zeros((4,), dtype=torch.int64, device=dev)
        sims[0].buf["explore_counters"] = cnt
        with torch.cuda.stream(stream):
            sims[0].explore_step()
        stream.synchronize()
        sims[0].buf["explore_counters"] = None
        explore_counts = [int(x) for x in cnt.cpu()]
        # EDF samples x 8 B + visible cells of the update x 16 B (RMW) + visible cells of the 9
        # candidates x 8 B, from the instrumented counters of one launch (all 10 poses share counter 2)
        xbytes = explore_counts[1] * 8 + explore_counts[2] * 8 + (explore_counts[2] // 10) * 8
        kern["k_explore"] = (t, xbytes, "timed alone", explore_counts[2] // 10 * 8 + 64 * NA // 2)
        t_sens += t
    if wl == "rvo":
        t = time_alone(lambda sm: sm.rvo_step())
        # per agent: own 3 state planes read (48 B) + 8 B action written; the neighbours' planes are re-read from
        # L1/L2 within the environment (A-1 x 48 B per agent, not HBM traffic)
        kern["k_rvo"] = (t, (48 + 8) * NA, "timed alone", (48 + 8) * NA)
        t_sens += t
    if wl == "c5":
        t_both = time_alone(lambda sm: sm.ga3c_forward(use_tensor_cores=use_tc, exact=ga3c_exact))
        t_fwd = time_alone(lambda sm: sm.ga3c_forward(use_tensor_cores=use_tc, forward_only=True))
        fwd_name = "k_ga3c_tc" if use_tc else "k_ga3c_fwd"
        kern[fwd_name] = (t_fwd, GA3C_FLOP_PER_AGENT * NA, "timed alone (forward kernel on the previous state vectors)", 0)
        # state-vector kernel: reads 4 state planes of the agent + 3 of each other agent (L1/L2-resident within the
        # env), writes the 76-float row + 11 fp16 chunks: counted as 64 + 304 + 176 B per agent
        t_plain = time_alone(lambda sm: sm.ga3c_forward(use_tensor_cores=use_tc)) if ga3c_exact else t_both
        kern["k_ga3c_state"] = (max(t_plain - t_fwd, 1e-9), (64 + 304 + 176) * NA, "state+forward timed alone minus the forward alone", (64 + 304 + 176) * NA)
        if ga3c_exact:
            kern["k_ga3c_fwd(fix-up)"] = (max(t_both - t_plain, 1e-9), 0, "exact mode minus the bare tensor-core forward: fp32 re-evaluation of the "
                                          "agents whose top-2 margin is below the fp16-operand error bound", 0)
        t_sens += t_both
    core_units = (BYTES_CORE + (BYTES_OAS if "other_agents_states" in W["sensors"] else 0)
                  + (BYTES_WALL if "laserscan" in W["sensors"] or "local_grid" in W["sensors"] else 0)) * NA
    core_must = (MUST_CORE + (MUST_OAS if "other_agents_states" in W["sensors"] else 0)) * NA
    kern["k_core"] = (max(step_s - t_sens, 1e-9), core_units, "timed alone" if not kern else "step time minus the kernels timed alone", core_must)
    dominant = max(kern, key=lambda k: kern[k][0])

    if rank == 0:
        cfgd = workload_config(wl, E, A, world, per_env_maps)
        line = {
            "metric": "agent-steps/s", "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": K, "warmup": Wm,
            "reps": reps, "timed_region_ms": ms, "ms_per_step": ms / (K * reps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfgd,
            "run": {"ring_batches": R, "resident_MB_per_gpu": R * batch_bytes / 1e6,
                    "launch": "one CUDA graph of K steps" if graph is not None else "one C-ABI call per step from Python",
                    "collective_in_timed_region": ("all_reduce(SUM) of the [8] statistics row of every step, side stream, "
                                                   + ("NCCL over %d ranks" % world if world > 1 else "1 rank: stream fork/join only"))
                    if collective else "off (--no-collective / --no-stats)",
                    "steps_per_launch": spl,
                    "ga3c_forward": args.ga3c if wl == "c5" else None,
                    "ga3c_weights": (args.ga3c_weights if ga3c_trained_weights() is not None else "random") if wl == "c5" else None},
            "e2e": {"value": e2e_value, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": Ke, "call": e2e_call},
            "gpu_launches": int(round(launches_per_step * K * reps)),
            "roofline": roofline_obj(wl, kern, dominant, step_s,
                                     sum(v[1] for k, v in kern.items() if not k.startswith("k_ga3c")),
                                     sum(v[3] for k, v in kern.items() if not k.startswith("k_ga3c")), explore_counts),
            "clocks": dict(sampler.summary(t_load0, t_load1), sampled_over="timed region + 1.2 s of the same workload"),
            "episode_stats": episode,
        }
        if not args.no_cpu_baseline:
            threads = min(os.cpu_count() or 1, 64)
            Es = min(REF_SAMPLE_ENVS[wl], E)
            n_steps = 100 if wl == "c2" else 20
            v, sps = oracle_run(wl, Es, A, n_steps, 2, threads, per_env_maps=per_env_maps)
            line["cpu_baseline"] = cpu_baseline_obj(wl, v, threads, "%d steps of %d envs x %d agents, oracle/casim_oracle.c with %d POSIX threads"
                                                    % (n_steps, Es, A, threads))
        emit(line)
    torch.cuda.synchronize()
    if world > 1:
        # Captured NCCL work keeps the communicator busy: destroy_process_group() / interpreter teardown with the graphs
        # still alive was seen to hang after the line had been printed (r2j, 2 ranks).  Every collective of this run has
        # completed (synchronize above, nothing follows the JSON line), so the ranks leave without the teardown.
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)
    for s in sims:
        s.close()


if __name__ == "__main__":
    main()
