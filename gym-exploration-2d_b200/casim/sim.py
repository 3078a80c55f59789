"""BatchedSim -- the device-resident state of E environments x A agent slots and thin wrappers around
the C ABI calls.  PyTorch is only the buffer / stream / interop layer: every computation is a kernel of
libcasim.so.

Layout (include/casim.h): ``state[6, E, A, 4]`` float32 (six float4 planes; step_num / flags / type are
int32 bit patterns), observations as dense ``[E, A, ...]`` tensors.
"""
import ctypes

import numpy as np
import torch

from . import layout as L
from . import lib as _lib
from .config import Config, SimConfig, make_sim_config, MAP_DIM, NUM_STATS


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


class BatchedSim:
    def __init__(self, num_envs, max_agents, cfg=None, device=None, sensors=(), want_debug=True, want_stats=True,
                 want_misc=True, stats_buffer=None):
        """``cfg``: a SimConfig (default: snapshot of casim.config.Config).  ``sensors``: iterable of sensor
        bit masks / names for which observation buffers are allocated ("other_agents_states", "laserscan",
        "local_grid")."""
        if not torch.cuda.is_available():
            raise _lib.CasimError("casim needs a CUDA device (B200); there is no CPU fallback")
        self.lib = _lib.load()
        self.E, self.A = int(num_envs), int(max_agents)
        self.cfg = cfg if cfg is not None else make_sim_config(Config)
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        h = ctypes.c_void_p()
        _lib.check(self.lib.casim_create(ctypes.byref(self.cfg), self.E, self.A, ctypes.byref(h)), "casim_create")
        self.handle = h
        E, A, dev = self.E, self.A, self.device
        f32 = dict(dtype=torch.float32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        self.state = torch.zeros((L.NUM_PACKS, E, A, 4), **f32)
        names = set()
        for s in sensors:
            if isinstance(s, str):
                names.add(s)
            else:
                if s & L.SENSOR_OTHER_AGENTS:
                    names.add("other_agents_states")
                if s & L.SENSOR_LASERSCAN:
                    names.add("laserscan")
                if s & L.SENSOR_OCCUPANCY_GRID:
                    names.add("local_grid")
        rows = self.cfg.max_other_agents_rows
        self.buf = {
            "obs_misc": torch.zeros((E, A, 4), **f32) if want_misc else None,
            "other_agents_states": torch.zeros((E, A, rows, 10), **f32) if "other_agents_states" in names else None,
            "laserscan": torch.zeros((E, A, self.cfg.laser_num_beams), **f32) if "laserscan" in names else None,
            "local_grid": torch.zeros((E, A, self.cfg.submap_w, self.cfg.submap_h), **u8) if "local_grid" in names else None,
            "rewards": torch.zeros((E, A), **f32),
            "done": torch.zeros((E, A), **u8),
            "game_over": torch.zeros((E,), **u8),
            "collision_flags": torch.zeros((E, A), **u8) if want_debug else None,
            "nearest_dist": torch.zeros((E, A), **f32) if want_debug else None,
            "stats": (stats_buffer if stats_buffer is not None else torch.zeros((NUM_STATS,), **f32)) if want_stats else None,
        }
        self._obs = self._obs_struct(self.buf)
        self._keep = {}          # tensors the library borrows (init, types, n_agents, traj)
        self.has_map = False

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _obs_struct(bufs):
        o = _lib.Obs()
        for name in _lib.OBS_FIELDS:
            t = bufs.get(name)
            setattr(o, name, None if t is None else _lib.ptr(t))
        return o

    def close(self):
        if getattr(self, "handle", None) is not None and self.handle:
            torch.cuda.synchronize(self.device)
            self.lib.casim_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self, a, dtype):
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            return a.to(device=self.device, dtype=dtype).contiguous()
        return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(self.device).contiguous()

    # ------------------------------------------------------------------------------------------
    def set_config(self, cfg):
        self.cfg = cfg
        _lib.check(self.lib.casim_set_config(self.handle, ctypes.byref(cfg)), "casim_set_config")

    def set_obstacles(self, rects, n_rects=None, want_map=False):
        """rects: [n_maps, max_rects, 4] (x1, y1, x3, y3) in metres, or [max_rects, 4] for one shared map,
        or None / empty for "no obstacles"."""
        if rects is None or len(rects) == 0:
            _lib.check(self.lib.casim_set_obstacles(self.handle, None, None, 0, 0, None, _stream_ptr()), "casim_set_obstacles")
            self.has_map = False
            return None
        r = self._dev(rects, torch.float32)
        if r.dim() == 2:
            r = r.unsqueeze(0)
        n_maps, max_rects = r.shape[0], r.shape[1]
        if n_rects is None:
            n_rects = torch.full((n_maps,), max_rects, dtype=torch.int32, device=self.device)
        else:
            n_rects = self._dev(n_rects, torch.int32)
        out = torch.zeros((n_maps, MAP_DIM, MAP_DIM), dtype=torch.uint8, device=self.device) if want_map else None
        _lib.check(self.lib.casim_set_obstacles(self.handle, _lib.ptr(r), _lib.ptr(n_rects), n_maps, max_rects,
                                                _lib.ptr(out), _stream_ptr()), "casim_set_obstacles")
        torch.cuda.current_stream().synchronize()      # r / n_rects may be freed after return
        self.has_map = True
        return out

    def set_map(self, static_map):
        """static_map: [300,300] or [n_maps,300,300] bool/uint8 (Map.static_map)."""
        if static_map is None:
            _lib.check(self.lib.casim_set_map(self.handle, None, 0, _stream_ptr()), "casim_set_map")
            self.has_map = False
            return
        m = self._dev(static_map, torch.uint8)
        if m.dim() == 2:
            m = m.unsqueeze(0)
        _lib.check(self.lib.casim_set_map(self.handle, _lib.ptr(m), m.shape[0], _stream_ptr()), "casim_set_map")
        torch.cuda.current_stream().synchronize()
        self.has_map = True

    def set_stats_event(self, event):
        """casim_set_stats_event: ``event`` (a torch.cuda.Event that has been recorded at least once, or None) is recorded
        by every ``step`` right after the fused step kernel, when the episode statistics are final."""
        self._keep["stats_event"] = event
        _lib.check(self.lib.casim_set_stats_event(self.handle, ctypes.c_void_p(event.cuda_event) if event is not None else None),
                   "casim_set_stats_event")

    def enable_packed_grid(self, on=True):
        """Second delivery format of ``local_grid`` (casim_set_grid_packed_output): ``buf['local_grid_packed']``
        uint8 [E, A, w*h/8] = numpy.packbits of every agent's grid, written by every later step next to the dense form."""
        if not on:
            _lib.check(self.lib.casim_set_grid_packed_output(self.handle, None), "casim_set_grid_packed_output")
            self.buf["local_grid_packed"] = None
            return None
        cells = self.cfg.submap_w * self.cfg.submap_h
        if self.cfg.submap_w % 4 or self.cfg.submap_h % 4:
            raise _lib.CasimError("packed local_grid needs submap sides that are multiples of 4")
        t = torch.zeros((self.E, self.A, cells // 8), dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.casim_set_grid_packed_output(self.handle, _lib.ptr(t)), "casim_set_grid_packed_output")
        self.buf["local_grid_packed"] = t
        return t

    def set_trajectories(self, traj):
        """traj: [E,A,T,4] (pose_x, pose_y, vel_x, vel_y) for ExternalDynamics agents."""
        if traj is None:
            self._keep["traj"] = None
            _lib.check(self.lib.casim_set_trajectories(self.handle, None, 0), "casim_set_trajectories")
            return
        t = self._dev(traj, torch.float32)
        self._keep["traj"] = t
        _lib.check(self.lib.casim_set_trajectories(self.handle, _lib.ptr(t), t.shape[2]), "casim_set_trajectories")

    def generate_scenario(self, kind, number_of_agents, seeds, ego_policy, ego_dynamics, other_dynamics=L.DYN_UNICYCLE,
                          out=None):
        """The reference's seeded scenario generators on the device (test_cases.py:1192-1463): kind =
        layout.SCENARIO_SWAP_CIRCLE / SCENARIO_RANDOM_POSITIONS, seeds [E] (non-zero).  Returns device tensors
        (init [E,A,7], types [E,A], n_agents [E], coop [E,A]) ready for ``reset``; ``out`` = such a tuple to overwrite
        in place (the generator is a pure function of the seed: rows whose seed did not change keep their bytes)."""
        seeds = torch.as_tensor(np.ascontiguousarray(seeds, dtype=np.uint32).view(np.int32)).to(self.device) \
            if not isinstance(seeds, torch.Tensor) else seeds.to(device=self.device, dtype=torch.int32).contiguous()
        assert seeds.shape == (self.E,)
        if out is not None:
            init, types, n, coop = out
        else:
            init = torch.empty((self.E, self.A, 7), dtype=torch.float32, device=self.device)
            types = torch.empty((self.E, self.A), dtype=torch.int32, device=self.device)
            n = torch.empty((self.E,), dtype=torch.int32, device=self.device)
            coop = torch.empty((self.E, self.A), dtype=torch.float32, device=self.device)
        _lib.check(self.lib.casim_generate_scenario(self.handle, int(kind), int(number_of_agents), _lib.ptr(seeds),
                                                    int(ego_policy), int(ego_dynamics), int(other_dynamics), _lib.ptr(init),
                                                    _lib.ptr(types), _lib.ptr(n), _lib.ptr(coop), _stream_ptr()),
                   "casim_generate_scenario")
        return init, types, n, coop

    def reset(self, init, types, n_agents=None):
        """init [E,A,7] = start_x, start_y, goal_x, goal_y, radius, pref_speed, heading (NaN = towards goal);
        types [E,A] int32 type words (layout.encode_type)."""
        init = self._dev(init, torch.float32)
        types = self._dev(types, torch.int32)
        n_agents = self._dev(n_agents, torch.int32)
        assert init.shape == (self.E, self.A, 7) and types.shape == (self.E, self.A)
        self._keep.update(init=init, types=types, n_agents=n_agents)
        _lib.check(self.lib.casim_reset(self.handle, _lib.ptr(self.state), _lib.ptr(init), _lib.ptr(types),
                                        _lib.ptr(n_agents), ctypes.byref(self._obs), _stream_ptr()), "casim_reset")

    def reset_envs(self, mask):
        """Restart the environments with mask[e] != 0 (uint8 / bool CUDA tensor [E], e.g. ``buf['game_over']``) from the
        init / types / n_agents tensors of the last ``reset`` (``_keep``; rewrite their rows first for a new scenario)
        and recompute their observations; every other environment is left untouched (casim_reset_envs)."""
        m = mask if mask.dtype == torch.uint8 else mask.to(torch.uint8)
        m = m.contiguous()
        _lib.check(self.lib.casim_reset_envs(self.handle, _lib.ptr(self.state), _lib.ptr(m), ctypes.byref(self._obs),
                                             _stream_ptr()), "casim_reset_envs")
        if getattr(self, "belief", None) is not None:            # a new episode starts from the prior target map
            self.belief[m.bool()] = float(self.cfg.ig_prior)
        for k in ("team_targets", "team_n"):                       # obsvd_targets is None before the first update
            for t in self.buf.get(k, []) or []:
                t[m.bool()] = 0

    def bind(self, types=None, n_agents=None, init=None):
        """Tell the library about n_agents / types without re-initialising the state (used by tests that
        inject a mid-episode state)."""
        init = self._dev(init if init is not None else np.zeros((self.E, self.A, 7), np.float32), torch.float32)
        types = self._dev(types, torch.int32)
        n_agents = self._dev(n_agents, torch.int32)
        scratch = torch.zeros_like(self.state)
        self._keep.update(init=init, types=types, n_agents=n_agents)
        o = self._obs_struct({})
        _lib.check(self.lib.casim_reset(self.handle, _lib.ptr(scratch), _lib.ptr(init), _lib.ptr(types),
                                        _lib.ptr(n_agents), ctypes.byref(o), _stream_ptr()), "casim_reset")
        torch.cuda.current_stream().synchronize()

    def observe(self):
        _lib.check(self.lib.casim_observe(self.handle, _lib.ptr(self.state), ctypes.byref(self._obs), _stream_ptr()),
                   "casim_observe")

    def step(self, ext_actions=None, policy_actions=None, dt=None):
        """One env.step for every environment, asynchronous on the current CUDA stream.
        ext_actions / policy_actions: float32 CUDA tensors [E,A,2] (or None)."""
        _lib.check(self.lib.casim_step(self.handle, _lib.ptr(self.state), _lib.ptr(ext_actions),
                                       _lib.ptr(policy_actions), float(dt) if dt is not None else 0.0,
                                       ctypes.byref(self._obs), _stream_ptr()), "casim_step")

    def step_n(self, n_steps, policy_actions=None, dt=None):
        """``n_steps`` env steps in ONE launch (casim_step_n; in-kernel policies only).  Returns the ring buffers
        dict(rewards [n,E,A], done [n,E,A], game_over [n,E], stats [n,8]); ``buf`` / ``state`` hold the observations and
        state after the last step (``buf['rewards']`` etc. are NOT updated by this call)."""
        n_steps = int(n_steps)
        ring = getattr(self, "_ring", None)
        if ring is None or ring["rewards"].shape[0] != n_steps:
            dev = self.device
            ring = dict(rewards=torch.zeros((n_steps, self.E, self.A), dtype=torch.float32, device=dev),
                        done=torch.zeros((n_steps, self.E, self.A), dtype=torch.uint8, device=dev),
                        game_over=torch.zeros((n_steps, self.E), dtype=torch.uint8, device=dev),
                        stats=torch.zeros((n_steps, NUM_STATS), dtype=torch.float32, device=dev) if self.buf["stats"] is not None else None)
            self._ring = ring
            self._ring_obs = self._obs_struct(dict(self.buf, **ring))
        _lib.check(self.lib.casim_step_n(self.handle, _lib.ptr(self.state), _lib.ptr(policy_actions),
                                         float(dt) if dt is not None else 0.0, n_steps, ctypes.byref(self._ring_obs),
                                         _stream_ptr()), "casim_step_n")
        return ring

    def run_sensor(self, which):
        """Launch one sensor kernel alone (profiling aid): which = layout.SENSOR_LASERSCAN / SENSOR_OCCUPANCY_GRID."""
        _lib.check(self.lib.casim_run_sensor(self.handle, _lib.ptr(self.state), ctypes.byref(self._obs), int(which),
                                             _stream_ptr()), "casim_run_sensor")

    class HostView:
        """Pre-built casim_obs over a fixed set of host tensors (keeps them alive); avoids rebuilding the struct on
        every ``step_host`` call."""

        def __init__(self, bufs):
            self.bufs = dict(bufs)
            self.struct = BatchedSim._obs_struct(self.bufs)

    def step_host(self, host_bufs, ext_actions_host=None, ext_actions_dev=None, policy_actions=None, dt=None):
        """Host-buffer step: H2D of the actions, the step, delivery of every buffer present in ``host_bufs``
        (dict name -> host tensor, or a ``BatchedSim.HostView`` of such a dict); returns when the results are in
        host memory.  Pinned scalar buffers are written by the kernel directly (see casim_step_host in casim.h)."""
        ho = host_bufs.struct if isinstance(host_bufs, BatchedSim.HostView) else self._obs_struct(host_bufs)
        _lib.check(self.lib.casim_step_host(self.handle, _lib.ptr(self.state), _lib.ptr(ext_actions_host),
                                            _lib.ptr(ext_actions_dev), _lib.ptr(policy_actions),
                                            float(dt) if dt is not None else 0.0, ctypes.byref(self._obs),
                                            ctypes.byref(ho), _stream_ptr()), "casim_step_host")

    def bind_host_step(self, host_bufs, ext_actions_host=None, ext_actions_dev=None, policy_actions=None, dt=None,
                       stream=None):
        """Pre-bound ``step_host``: every pointer / struct / the stream handle is resolved ONCE and the returned
        callable is a bare ``casim_step_host`` call (an RL loop steps the same buffers every iteration; building the
        ctypes arguments costs more host time than the C-ABI call).  The tensors must stay alive and must not be
        re-allocated while the callable is in use; ``stream`` defaults to the stream current at bind time."""
        view = host_bufs if isinstance(host_bufs, BatchedSim.HostView) else BatchedSim.HostView(host_bufs)
        stream = torch.cuda.current_stream(self.device) if stream is None else stream
        fn, check = self.lib.casim_step_host, _lib.check
        args = (self.handle, _lib.ptr(self.state), _lib.ptr(ext_actions_host), _lib.ptr(ext_actions_dev),
                _lib.ptr(policy_actions), float(dt) if dt is not None else 0.0, ctypes.byref(self._obs),
                ctypes.byref(view.struct), ctypes.c_void_p(stream.cuda_stream))
        keep = (view, ext_actions_host, ext_actions_dev, policy_actions, stream)

        def call(_keep=keep):
            rc = fn(*args)
            if rc:
                check(rc, "casim_step_host")
        return call

    def host_pipeline(self, fields, packed_grid=True, ext_actions=False, stream=None):
        """Pipelined host delivery (casim_step_host_pipelined): returns a ``HostPipeline`` whose ``step()`` enqueues one
        env step and hands back the PREVIOUS step's results in pinned host tensors (one step of latency, D2H copies
        overlapped with the next step's kernels).  ``fields``: names of the result / observation buffers wanted on the
        host; with ``packed_grid`` the occupancy grid travels as numpy.packbits bytes (``'local_grid_packed'``)."""
        return HostPipeline(self, fields, packed_grid, ext_actions, stream)

    # ------------------------------------------------------------------------------------------
    # exploration (ig_greedy) and GA3C-CADRL policies: both write into the shared ``policy_actions`` buffer
    def _policy_buffers(self):
        if "policy_actions" not in self.buf or self.buf.get("policy_actions") is None:
            self.buf["policy_actions"] = torch.zeros((self.E, self.A, 2), dtype=torch.float32, device=self.device)
        return self.buf["policy_actions"]

    def enable_exploration(self, want_debug=True):
        """Allocate the target-map belief [E,A,dim,dim] (fp32 odds, prior everywhere) and the result buffers of
        casim_explore_step."""
        dim = self.cfg.ig_map_dim
        f32 = dict(dtype=torch.float32, device=self.device)
        self.belief = torch.full((self.E, self.A, dim, dim), float(self.cfg.ig_prior), **f32)
        self._policy_buffers()
        self.buf["best_mi"] = torch.zeros((self.E, self.A), **f32) if want_debug else None
        self.buf["cand_mi"] = torch.zeros((self.E, self.A, 9), **f32) if want_debug else None
        self.buf["explore_counters"] = torch.zeros((4,), dtype=torch.int64, device=self.device) if want_debug else None

    def explore_step(self):
        """ig_greedy.find_next_action for every ig_greedy agent: belief update + greedy action into
        ``buf['policy_actions']`` (pass it to ``step(policy_actions=...)``)."""
        if self.buf.get("other_agents_states") is None:
            raise _lib.CasimError("explore_step needs the other_agents_states observation buffer")
        _lib.check(self.lib.casim_explore_step(self.handle, _lib.ptr(self.state), _lib.ptr(self.buf["other_agents_states"]),
                                               _lib.ptr(self.belief), _lib.ptr(self.buf["policy_actions"]),
                                               _lib.ptr(self.buf.get("best_mi")), _lib.ptr(self.buf.get("cand_mi")),
                                               _lib.ptr(self.buf.get("explore_counters")), _stream_ptr()),
                   "casim_explore_step")
        return self.buf["policy_actions"]

    def explore_team_update(self):
        """ig_mcts.update_belief for every explorer agent: the belief is updated from the agent's own pose and
        detections and from the other explorers' poses and detection lists (``buf['team_targets']`` ping-pong)."""
        if getattr(self, "belief", None) is None:
            raise _lib.CasimError("explore_team_update needs enable_exploration() first")
        if self.buf.get("other_agents_states") is None:
            raise _lib.CasimError("explore_team_update needs the other_agents_states observation buffer")
        if "team_targets" not in self.buf:
            dev = self.device
            self.buf["team_targets"] = [torch.zeros((self.E, self.A, 32, 2), dtype=torch.float64, device=dev) for _ in range(2)]
            self.buf["team_n"] = [torch.zeros((self.E, self.A), dtype=torch.int32, device=dev) for _ in range(2)]
            self.buf["team_mask"] = torch.zeros((self.E, self.A, 128), dtype=torch.int32, device=dev)
            self.buf["team_reward"] = torch.zeros((self.E, self.A), dtype=torch.float32, device=dev)
        (tp, tn), (cp, cn) = self.buf["team_targets"], self.buf["team_n"]
        _lib.check(self.lib.casim_explore_team_update(self.handle, _lib.ptr(self.state), _lib.ptr(self.buf["other_agents_states"]),
                                                      _lib.ptr(self.belief), _lib.ptr(tp), _lib.ptr(cp), _lib.ptr(tn),
                                                      _lib.ptr(cn), _lib.ptr(self.buf["team_mask"]),
                                                      _lib.ptr(self.buf["team_reward"]), _stream_ptr()),
                   "casim_explore_team_update")
        self.buf["team_targets"], self.buf["team_n"] = [tn, tp], [cn, cp]      # index 0 = the latest lists
        return self.buf["team_mask"], self.buf["team_reward"]

    def reset_team_targets(self):
        """Forget the detection lists (a new episode: obsvd_targets is None before the first update, ig_mcts.py:127)."""
        for k in ("team_targets", "team_n"):
            for t in self.buf.get(k, []):
                t.zero_()

    def explore_rollouts(self, actions, exclude=None, xdt=1, want_end_pose=True, want_obsvd=True):
        """Dec-MCTS simulation rollouts (ig_mcts.get_next_pose / mcts_sim_state_storer / mcts_reward, the body of
        pydecmcts Tree._simulate) for every explorer agent from its current pose.
        actions: uint8 [E,A,K,H] (3*vel_idx + dphi_idx, > 8 ends a sequence); exclude: packed cell masks
        (``pack_cell_mask``) [E,A,128] or [E,A,K,128] of the other robots' cells, or None.
        Returns dict(rewards [E,A,K], end_pose [E,A,K,3] | None, obsvd [E,A,K,128] int32 | None,
        n_feasible [E,A,K] int32 = steps of each sequence that were applied)."""
        if getattr(self, "belief", None) is None:
            raise _lib.CasimError("explore_rollouts needs enable_exploration() first")
        actions = torch.as_tensor(actions, dtype=torch.uint8, device=self.device).contiguous()
        if actions.dim() != 4 or tuple(actions.shape[:2]) != (self.E, self.A):
            raise _lib.CasimError("actions must be [E, A, K, H]")
        K, H = int(actions.shape[2]), int(actions.shape[3])
        per = 0
        if exclude is not None:
            exclude = torch.as_tensor(exclude, dtype=torch.int32, device=self.device).contiguous()
            if tuple(exclude.shape) == (self.E, self.A, K, 128):
                per = 1
            elif tuple(exclude.shape) != (self.E, self.A, 128):
                raise _lib.CasimError("exclude must be [E, A, 128] or [E, A, K, 128]")
        f32 = dict(dtype=torch.float32, device=self.device)
        rewards = torch.empty((self.E, self.A, K), **f32)
        end_pose = torch.empty((self.E, self.A, K, 3), **f32) if want_end_pose else None
        obsvd = torch.empty((self.E, self.A, K, 128), dtype=torch.int32, device=self.device) if want_obsvd else None
        n_feasible = torch.empty((self.E, self.A, K), dtype=torch.int32, device=self.device)
        _lib.check(self.lib.casim_explore_rollouts(self.handle, _lib.ptr(self.state), _lib.ptr(self.belief),
                                                   _lib.ptr(actions), K, H, int(xdt), _lib.ptr(exclude), per,
                                                   _lib.ptr(rewards), _lib.ptr(end_pose), _lib.ptr(obsvd), _lib.ptr(n_feasible),
                                                   _stream_ptr()),
                   "casim_explore_rollouts")
        return dict(rewards=rewards, end_pose=end_pose, obsvd=obsvd, n_feasible=n_feasible)

    def ga3c_set_weights(self, weights, want_debug=True):
        """weights: float32 [170764] parameter blob in the order of include/casim.h (TF variables of
        GA3C_CADRL/network.py: lstm kernel/bias, layer1, layer2, fullyconnected1, logits_p, logits_v)."""
        w = np.ascontiguousarray(weights, dtype=np.float32)
        _lib.check(self.lib.casim_ga3c_set_weights(self.handle, _lib.ptr(w), int(w.size), _stream_ptr()),
                   "casim_ga3c_set_weights")
        self._policy_buffers()
        dev = self.device
        self.buf["action_idx"] = torch.zeros((self.E, self.A), dtype=torch.int32, device=dev) if want_debug else None
        self.buf["probs"] = torch.zeros((self.E, self.A, 11), dtype=torch.float32, device=dev) if want_debug else None
        self.buf["nn_input"] = torch.zeros((self.E, self.A, 75), dtype=torch.float32, device=dev) if want_debug else None

    def ga3c_forward(self, use_tensor_cores=False, forward_only=False, exact=False):
        """GA3CCADRLPolicy.find_next_action for every GA3C agent -> ``buf['policy_actions']``.
        ``use_tensor_cores``: tcgen05 forward (fp16 operands, fp32 accumulate); with ``exact`` the agents it cannot call (top-2 margin below the fp16
        error bound) are re-evaluated in fp32 in the same call, so the actions equal the fp32 forward's.
        ``forward_only`` (profiling aid) skips the state-vector kernel and reuses the previous call's vectors."""
        _lib.check(self.lib.casim_ga3c_forward(self.handle, _lib.ptr(self.state), _lib.ptr(self.buf["policy_actions"]),
                                               _lib.ptr(self.buf.get("action_idx")), _lib.ptr(self.buf.get("probs")),
                                               _lib.ptr(self.buf.get("nn_input")),
                                               int(bool(use_tensor_cores)) | (2 if forward_only else 0) | (4 if (exact and use_tensor_cores) else 0),
                                               _stream_ptr()), "casim_ga3c_forward")
        return self.buf["policy_actions"]

    def rvo_set_obstacles(self, polygons):
        """sim.addObstacle for every polygon + sim.processObstacles (RVOPolicy.py:44,56-57): ``polygons`` = list of
        vertex lists [(x, y), ...] (counter-clockwise), ONE set shared by all environments; None / [] removes it."""
        polygons = [] if polygons is None else [np.asarray(p, np.float64).reshape(-1, 2) for p in polygons]
        sizes = (ctypes.c_int32 * max(len(polygons), 1))(*[len(p) for p in polygons])
        flat = np.ascontiguousarray(np.concatenate(polygons) if polygons else np.zeros((0, 2)), np.float64)
        _lib.check(self.lib.casim_rvo_set_obstacles(self.handle, flat.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), sizes,
                                                    len(polygons)), "casim_rvo_set_obstacles")

    def rvo_step(self, coop=None, params=None):
        """RVOPolicy.find_next_action for every RVO agent -> ``buf['policy_actions']`` (pass it to
        ``step(policy_actions=...)``).  coop: [E,A] float32 Agent.cooperation_coef (None = 1.0)."""
        self._policy_buffers()
        if coop is not None:
            coop = self._dev(coop, torch.float32)
            self._keep["rvo_coop"] = coop
        _lib.check(self.lib.casim_rvo_step(self.handle, _lib.ptr(self.state), _lib.ptr(coop),
                                           ctypes.byref(params) if params is not None else None,
                                           _lib.ptr(self.buf["policy_actions"]), _stream_ptr()), "casim_rvo_step")
        return self.buf["policy_actions"]

    def flat_row_size(self, keys):
        ids = (ctypes.c_int32 * len(keys))(*[L.OBS_KEY_ID[k] for k in keys])
        n = int(self.lib.casim_flat_row_size(self.handle, ids, len(keys)))
        if n < 0:
            raise _lib.CasimError("unknown observation key in %r" % (keys,))
        return n

    def flatten_obs(self, keys, out=None):
        """MultiagentFlattenDictWrapper.observation (envs/wrappers.py:37-45) for every environment: float32
        [E, A*S] with, per agent slot, the listed keys ravel()ed in order (absent slots zero)."""
        ids = (ctypes.c_int32 * len(keys))(*[L.OBS_KEY_ID[k] for k in keys])
        S = self.flat_row_size(keys)
        if out is None:
            out = torch.empty((self.E, self.A * S), dtype=torch.float32, device=self.device)
        _lib.check(self.lib.casim_flatten_obs(self.handle, _lib.ptr(self.state), ctypes.byref(self._obs), ids, len(keys),
                                              _lib.ptr(out), _stream_ptr()), "casim_flatten_obs")
        return out

    def launch_count(self):
        return int(self.lib.casim_launch_count(self.handle))

    # ------------------------------------------------------------------------------------------
    # host-side views (tests / drop-in shim)
    def state_numpy(self):
        """float64 [E,A,24] copy of the state (int words decoded)."""
        packs = self.state.detach().cpu().numpy()
        P, E, A, _ = packs.shape
        w = np.ascontiguousarray(packs.transpose(1, 2, 0, 3).reshape(E, A, P * 4))
        out = w.astype(np.float64)
        iw = w.view(np.int32)
        for k in L.INT_WORDS:
            out[..., k] = iw[..., k]
        return out

    def load_state_numpy(self, state):
        """Inverse of state_numpy: float64 [E,A,24] -> device state (values rounded to fp32)."""
        E, A, W = state.shape
        w = state.astype(np.float32)
        iw = w.view(np.int32)
        for k in L.INT_WORDS:
            iw[..., k] = state[..., k].astype(np.int32)
        packs = np.ascontiguousarray(w.reshape(E, A, W // 4, 4).transpose(2, 0, 1, 3))
        self.state.copy_(torch.from_numpy(packs).to(self.device))


class HostPipeline:
    """Two alternating sets of device observation buffers + pinned host buffers around casim_step_host_pipelined.
    ``step()`` returns the dict of host tensors holding the previous step's results (None on the very first call);
    ``flush()`` returns the last step's.  A returned dict is valid until the second ``step()`` after it."""

    def __init__(self, sim, fields, packed_grid, ext_actions, stream):
        self.sim = sim
        fields = list(fields)
        self.packed = bool(packed_grid and "local_grid" in fields)
        if self.packed:
            fields.remove("local_grid")
        dev_sets = [dict(sim.buf), {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in sim.buf.items()}]
        self.dev = [BatchedSim._obs_struct(d) for d in dev_sets]
        self._dev_keep = dev_sets
        self.host = []
        self.host_structs = []
        self.pk_dev, self.pk_host = [None, None], [None, None]
        cells = sim.cfg.submap_w * sim.cfg.submap_h
        for k in range(2):
            hb = {f: torch.zeros(sim.buf[f].shape, dtype=sim.buf[f].dtype).pin_memory() for f in fields}
            self.host_structs.append(BatchedSim._obs_struct(hb))
            if self.packed:
                self.pk_dev[k] = torch.zeros((sim.E, sim.A, cells // 8), dtype=torch.uint8, device=sim.device)
                self.pk_host[k] = torch.zeros((sim.E, sim.A, cells // 8), dtype=torch.uint8).pin_memory()
                hb = dict(hb, local_grid_packed=self.pk_host[k])
            self.host.append(hb)
        self.act_host = torch.zeros((sim.E, sim.A, 2), dtype=torch.float32).pin_memory() if ext_actions else None
        self.act_dev = torch.zeros((sim.E, sim.A, 2), dtype=torch.float32, device=sim.device) if ext_actions else None
        self.stream = torch.cuda.current_stream(sim.device) if stream is None else stream
        self.t = 0
        self.h2d_bytes = self.act_host.numel() * 4 if ext_actions else 0
        self.d2h_bytes = sum(v.numel() * v.element_size() for v in self.host[0].values())
        fn = sim.lib.casim_step_host_pipelined
        self._calls = []
        for k in range(2):
            self._calls.append((fn, (sim.handle, _lib.ptr(sim.state), _lib.ptr(self.act_host), _lib.ptr(self.act_dev)),
                                (ctypes.byref(self.dev[k]), ctypes.byref(self.host_structs[k]), _lib.ptr(self.pk_dev[k]),
                                 _lib.ptr(self.pk_host[k]), ctypes.c_void_p(self.stream.cuda_stream))))

    def step(self, policy_actions=None, dt=None):
        k = self.t & 1
        fn, head, tail = self._calls[k]
        rc = fn(*head, _lib.ptr(policy_actions), float(dt) if dt is not None else 0.0, *tail)
        if rc:
            _lib.check(rc, "casim_step_host_pipelined")
        self.t += 1
        return self.host[k ^ 1] if self.t > 1 else None

    def flush(self):
        _lib.check(self.sim.lib.casim_host_flush(self.sim.handle), "casim_host_flush")
        return self.host[(self.t - 1) & 1] if self.t > 0 else None


def pack_cell_mask(mask):
    """bool [..., dim, dim] (row j = y cell, column i = x cell, dim <= 64) -> int32 [..., 128] with bit
    (j * 64 + i) set for every true cell: the mask layout of casim_explore_rollouts (include/casim.h)."""
    m = np.asarray(mask, dtype=bool)
    dim = m.shape[-1]
    full = np.zeros(m.shape[:-2] + (64, 64), dtype=np.uint8)
    full[..., :dim, :dim] = m
    bits = np.packbits(full.reshape(m.shape[:-2] + (128, 32)), axis=-1, bitorder="little")
    return np.ascontiguousarray(bits).view(np.uint32).reshape(m.shape[:-2] + (128,)).view(np.int32)


def unpack_cell_mask(words, dim):
    """Inverse of ``pack_cell_mask``: int32/uint32 [..., 128] -> bool [..., dim, dim]."""
    w = np.ascontiguousarray(np.asarray(words)).view(np.uint32)
    b = np.unpackbits(w.view(np.uint8).reshape(w.shape[:-1] + (512,)), axis=-1, bitorder="little")
    return b.reshape(w.shape[:-1] + (64, 64))[..., :dim, :dim].astype(bool)
