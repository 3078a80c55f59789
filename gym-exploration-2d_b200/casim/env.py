"""BatchedCollisionAvoidanceEnv -- the gym-shaped boundary of the reference's CollisionAvoidanceEnv
(envs/collision_avoidance_env.py:53-232) for E environments at once.

Kept from the reference: ``reset() -> obs``, ``step(actions, dt=None) -> (obs, rewards, game_over,
{'which_agents_done': ...})``, ``set_agents``, ``set_static_map``, the observation names of
``Config.STATE_INFO_DICT`` (envs/config.py:104-215) and the ``Agent`` / ``Policy`` / ``Dynamics`` / ``Sensor`` plugin
vocabulary (casim.plugins).  Every observation entry gains leading ``[E, A]`` dims and is a CUDA tensor;
``obs[i][name]`` (E == 1) and ``obs[e][i][name]`` give the reference's per-agent view for drop-in use.
All computation happens in libcasim.so (casim.sim.BatchedSim); there is no CPU path.
"""
import numpy as np
import torch

from . import layout as L
from . import lib as _lib
from .config import Config, make_sim_config
from . import spaces as _spaces
from .plugins import Agent
from .sim import BatchedSim

# observation name -> (source, index): envs/config.py:104-215 'attr' strings, re-expressed on the SoA tensors
_STATE_WORDS = {"dist_to_goal": L.DIST_TO_GOAL, "radius": L.RADIUS, "heading_ego_frame": L.HEADING_EGO,
                "heading_global_frame": L.HEADING, "pref_speed": L.PREF_SPEED}


class _AgentObs:
    """obs[e][i]: dict-like view of one agent's observation (numpy, like the reference's values)."""

    def __init__(self, obs, e, i):
        self._obs, self._e, self._i = obs, e, i

    def __getitem__(self, name):
        return self._obs[name][self._e, self._i].detach().cpu().numpy()

    def keys(self):
        return self._obs.keys()

    def __contains__(self, name):
        return name in self._obs.keys()


class _EnvObs:
    def __init__(self, obs, e):
        self._obs, self._e = obs, e

    def __getitem__(self, i):
        return _AgentObs(self._obs, self._e, int(i))


class BatchedObs:
    """obs[name] -> CUDA tensor [E, A, ...]; obs[i] (E == 1) / obs[e][i] -> per-agent dict view."""

    def __init__(self, env):
        self._env = env

    def keys(self):
        return list(self._env.states_in_obs)

    def __getitem__(self, key):
        if isinstance(key, str):
            return self._env._obs_tensor(key)
        if self._env.E == 1:
            return _AgentObs(self, 0, int(key))
        return _EnvObs(self, int(key))


class BatchedCollisionAvoidanceEnv:
    def __init__(self, num_envs=1, max_agents=None, config=Config, device=None):
        self.E = int(num_envs)
        self.A = int(max_agents if max_agents is not None else config.MAX_NUM_AGENTS_IN_ENVIRONMENT)
        self.config = config
        self.cfg = make_sim_config(config)
        self.device = device
        self.dt_nominal = config.DT
        self.agents = None              # list (E == 1) or list of lists of casim.plugins.Agent
        self.obstacles = None
        self.sim = None
        self._tensors = None            # (init, types, n_agents) when set without Agent objects
        self._traj = None
        self._ga3c_weights = None
        self.states_in_obs = list(config.STATES_IN_OBS)
        # gym-facing attributes (collision_avoidance_env.py:97-134); per environment, like the reference's
        self.max_heading_change, self.min_heading_change = 4, -4
        self.min_speed, self.max_speed = -4, 4
        self.action_space_type = config.ACTION_SPACE_TYPE
        self.action_space = _spaces.action_space_from_env(self)
        self.observation_space = _spaces.observation_space_from_config(config)
        self.num_envs = self.E
        self.episode_number = 0
        self.episode_step_number = 0
        self._host_state = None
        self._generator = None
        # GA3C-CADRL forward: tensor cores (fp16 operands) with the fp32 fix-up of the agents they cannot call ("exact"): the action
        # index equals the fp32 forward's for every agent.  ga3c_exact = False gives the bare tensor-core forward,
        # ga3c_use_tensor_cores = False the fp32 forward.
        self.ga3c_use_tensor_cores = True
        self.ga3c_exact = True

    # ---- scenario injection (the reference's reset() ignores set_agents(): SURVEY.md 3.2) ----
    def set_agents(self, agents):
        """list[Agent] (shared by all environments) or one list per environment"""
        if len(agents) and isinstance(agents[0], Agent):
            agents = [agents] * self.E
        assert len(agents) == self.E
        self.agents = agents
        self._tensors = None

    def set_scenario_tensors(self, init, types, n_agents=None):
        """init [E,A,7], types [E,A] (layout.encode_type), n_agents [E]: the batched form of set_agents"""
        self._tensors = (np.asarray(init, np.float32), np.asarray(types, np.int32),
                         None if n_agents is None else np.asarray(n_agents, np.int32))
        self.agents = None

    def set_scenario_generator(self, kind, number_of_agents, ego_policy, ego_dynamics, other_dynamics=None, seed_base=0):
        """Generate every episode's agents on the device, like _init_agents in evaluate mode does on the host
        (collision_avoidance_env.py:403-415: ``tc.<scenario>(number_of_agents=..., seed=episode_number)``): kind =
        layout.SCENARIO_*, ego_policy / dynamics = plugin classes or ids.  Environment e of episode k uses seed
        ``seed_base + k * num_envs + e + 1`` (E == 1, seed_base == 0: the reference's seed = episode_number)."""
        ident = lambda x, attr: x if isinstance(x, int) else getattr(x, attr)
        od = L.DYN_UNICYCLE if other_dynamics is None else ident(other_dynamics, "dynamics_id")
        self._generator = (int(kind), int(number_of_agents), ident(ego_policy, "policy_id"),
                           ident(ego_dynamics, "dynamics_id"), od, int(seed_base))
        self.agents = None
        self._tensors = None

    def set_static_map(self, obstacles):
        """obstacles: the reference's format (list of 4-corner lists; corners 1 and 3 are read, Map.py:117-118), or
        an array [n,4] / [E,n,4] of (x1, y1, x3, y3) rectangles, or None."""
        self._obstacle_polys = None
        if obstacles is None or len(obstacles) == 0:
            self.obstacles = None
            return
        ob = obstacles
        if not isinstance(ob, np.ndarray) and hasattr(ob[0][0], "__len__"):
            self._obstacle_polys = [np.asarray(o, np.float64) for o in ob]     # what RVOPolicy hands to sim.addObstacle
        if not isinstance(ob, np.ndarray) and len(ob[0]) == 4 and hasattr(ob[0][0], "__len__"):
            ob = np.array([[o[1][0], o[1][1], o[3][0], o[3][1]] for o in ob], np.float32)
        self.obstacles = np.asarray(ob, np.float32)

    def set_trajectories(self, traj):
        self._traj = traj

    def initialize_network(self, checkpoints_root, **kwargs):
        """GA3CCADRLPolicy.initialize_network (GA3CCADRLPolicy.py:21-32): load a shipped TF1 checkpoint
        (kwargs checkpt_dir, checkpt_name as in the reference; checkpoints_root = .../GA3C_CADRL/checkpoints)."""
        from .ga3c_checkpoint import checkpoint_prefix, ga3c_weights_from_checkpoint
        self.set_ga3c_weights(ga3c_weights_from_checkpoint(checkpoint_prefix(checkpoints_root, **kwargs)))

    def set_ga3c_weights(self, weights):
        """float32 [170764] parameter blob (GA3CCADRLPolicy.initialize_network's counterpart)"""
        self._ga3c_weights = np.asarray(weights, np.float32)

    # ---- helpers ----
    def _scenario(self):
        if self._tensors is not None:
            return self._tensors
        if self.agents is None:
            # collision_avoidance_env.py:403-445 _init_agents: tc.<Config.SCENARIOS_FOR_TRAINING[...]>(); the fork's list
            # is ["IG_agent_crossing"] (config.py:91), a fixed scenario -> every environment of the batch starts it
            from . import test_cases as tc
            names = getattr(self.config, "SCENARIOS_FOR_TRAINING", None) or []
            made = tc.default_scenario(names[0]) if names else None
            if made is None:
                raise RuntimeError("call set_agents(), set_scenario_tensors() or set_scenario_generator() before reset() "
                                   "(Config.SCENARIOS_FOR_TRAINING names no fixed scenario)")
            if len(made[0]) > self.A:
                raise RuntimeError("scenario %s has %d agents, max_agents is %d" % (names[0], len(made[0]), self.A))
            self.set_agents([tc.default_scenario(names[0])[0] for _ in range(self.E)])
            self.set_static_map(made[1])
        init = np.zeros((self.E, self.A, 7), np.float32)
        types = np.zeros((self.E, self.A), np.int32)
        n = np.zeros((self.E,), np.int32)
        for e, ags in enumerate(self.agents):
            if len(ags) > self.A:
                raise ValueError("environment %d has %d agents, max_agents is %d" % (e, len(ags), self.A))
            n[e] = len(ags)
            for i, a in enumerate(ags):
                init[e, i] = a.init_row()
                types[e, i] = a.type_word
                a._env, a._e, a._i = self, e, i
        return init, types, n

    def _state_word(self, e, i, k):
        if self._host_state is None:
            self._host_state = self.sim.state_numpy()
        return float(self._host_state[e, i, k])

    def _obs_tensor(self, name):
        s, b = self.sim, self.sim.buf
        if name in _STATE_WORDS:
            k = _STATE_WORDS[name]
            return s.state[k // 4, :, :, k % 4]
        if name == "pos_global_frame":
            return s.state[0, :, :, 0:2]
        if name == "rel_goal":
            return b["obs_misc"][..., 0:2]
        if name == "num_other_agents":
            return b["obs_misc"][..., 2]
        if name == "use_ppo":
            return b["obs_misc"][..., 3]
        if name == "other_agent_states":        # config.py:171-177: the first row of other_agents_states (sic)
            return b["other_agents_states"][:, :, 0, :]
        if name in ("other_agents_states", "laserscan", "local_grid"):
            t = b.get(name)
            if t is None:
                raise KeyError("observation '%s' needs its sensor on at least one agent" % name)
            return t
        raise KeyError(name)

    # ---- gym API ----
    def reset(self):
        """collision_avoidance_env.py:234-266: (re)create the agents' state, build the static map, first _get_obs"""
        gen_coop = None
        if self._generator is not None:
            kind, n_ag, ego_pol, ego_dyn, other_dyn, base = self._generator
            if self.sim is None:    # the generators' sensor lists are fixed (test_cases.py:1262, :1279, :1445)
                bits = L.SENSOR_OTHER_AGENTS | (L.SENSOR_OCCUPANCY_GRID if kind == L.SCENARIO_SWAP_CIRCLE else 0)
                self.sim = BatchedSim(self.E, self.A, cfg=self.cfg, device=self.device, sensors=(bits,))
            self._env_episode = np.full((self.E,), self.episode_number, np.int64)
            seeds = (base + self._env_episode * self.E + np.arange(self.E) + 1).astype(np.uint32)
            self._gen_out = self.sim.generate_scenario(kind, n_ag, seeds, ego_pol, ego_dyn, other_dyn)
            init, types_dev, n, gen_coop = self._gen_out
            types = types_dev.cpu().numpy()
        else:
            init, types, n = self._scenario()
        sensor_bits = int(np.bitwise_or.reduce((types >> 16).ravel() & 0xff)) if types.size else 0
        policies = set((types & 0xff).ravel().tolist())
        if L.POLICY_IG_GREEDY in policies or L.POLICY_GA3C_CADRL in policies:
            sensor_bits |= L.SENSOR_OTHER_AGENTS
        if self.sim is None:
            self.sim = BatchedSim(self.E, self.A, cfg=self.cfg, device=self.device, sensors=(sensor_bits,))
        sim = self.sim
        sim.set_obstacles(self.obstacles)
        sim.set_trajectories(self._traj)
        self._has_ig = L.POLICY_IG_GREEDY in policies
        self._has_ga3c = L.POLICY_GA3C_CADRL in policies
        self._has_rvo = L.POLICY_RVO in policies
        self._rvo_coop = None
        if self._has_rvo:
            if self.obstacles is not None:
                if getattr(self, "_obstacle_polys", None) is None:
                    raise RuntimeError("RVOPolicy agents need the obstacles as polygons (the reference's list-of-corners "
                                       "format), not as rectangles")
                sim.rvo_set_obstacles(self._obstacle_polys)
            else:
                sim.rvo_set_obstacles(None)
            coop = np.ones((self.E, self.A), np.float32)
            if self.agents is not None:
                for e, ags in enumerate(self.agents):
                    for i, a in enumerate(ags):
                        coop[e, i] = a.cooperation_coef
            self._rvo_coop = torch.from_numpy(coop).to(sim.device) if gen_coop is None else gen_coop
            self._rvo_params = _lib.rvo_default_params()
            self._rvo_params.neighbor_dist = float(self.config.SENSING_HORIZON)
            self._rvo_params.time_horizon = float(getattr(self.config, "RVO_TIME_HORIZON", 5.0))
            self._rvo_params.max_neighbors = int(self.config.MAX_NUM_AGENTS_IN_ENVIRONMENT)
        self._needs_actions = bool(policies & {L.POLICY_EXTERNAL, L.POLICY_LEARNING, L.POLICY_CARRL})
        if self._has_ig:
            if self.obstacles is None:
                raise RuntimeError("ig_greedy agents need a static map (set_static_map)")
            sim.enable_exploration()
        self._mcts = None
        self._mcts_params = None
        if self._has_ig and self.agents is not None:
            pols = [a.policy for ags in self.agents for a in ags if getattr(a.policy, "str", "") == "ig_mcts"]
            if pols:                                   # the planner parameters are the first ig_mcts agent's (:344)
                self._mcts_params = pols[0]
        if self._has_ga3c:
            if self._ga3c_weights is None:
                raise RuntimeError("GA3CCADRLPolicy agents need set_ga3c_weights() (random-init or a loaded checkpoint)")
            sim.ga3c_set_weights(self._ga3c_weights)
        # generator path: hand the device tensors themselves to the library (it borrows them; reset_done_envs
        # regenerates their rows in place)
        sim.reset(init, self._gen_out[1] if self._generator is not None else types, n)
        self._types = types
        if self._has_ig and getattr(self, "_mcts_params", None) is not None:
            from .mcts import TeamPlanner
            sim.reset_team_targets()
            self._mcts = TeamPlanner(sim, self._mcts_params)
        self._ext = torch.zeros((self.E, self.A, 2), dtype=torch.float32, device=sim.device) if self._needs_actions else None
        # only the entries whose sensors exist are emitted (the reference yields np.array(None) for the others)
        avail = {"other_agents_states": sim.buf["other_agents_states"], "other_agent_states": sim.buf["other_agents_states"],
                 "laserscan": sim.buf["laserscan"], "local_grid": sim.buf["local_grid"]}
        self.states_in_obs = [s for s in self.config.STATES_IN_OBS if avail.get(s, True) is not None]
        self.episode_number += 1
        self.episode_step_number = 0
        self._host_state = None
        return BatchedObs(self)

    def _actions_tensor(self, actions):
        if actions is None or (isinstance(actions, dict) and not actions):
            return self._ext
        if isinstance(actions, torch.Tensor):
            return actions.to(device=self.sim.device, dtype=torch.float32).reshape(self.E, self.A, 2).contiguous()
        if isinstance(actions, dict):
            # reference form {agent_index: action} (E == 1) or {(env, agent_index): action}
            host = np.zeros((self.E, self.A, 2), np.float32)
            for k, v in actions.items():
                e, i = (0, k) if not isinstance(k, tuple) else k
                host[e, i, :len(np.atleast_1d(v))] = np.atleast_1d(v)
            self._ext.copy_(torch.from_numpy(host))
            return self._ext
        a = np.asarray(actions, np.float32).reshape(self.E, self.A, 2)
        self._ext.copy_(torch.from_numpy(a))
        return self._ext

    def step(self, actions=None, dt=None):
        """collision_avoidance_env.py:162-232.  Returns (obs, rewards, game_over, {'which_agents_done': done}) with
        rewards [E,A] (or [E] = agent 0's when Config.TRAIN_SINGLE_AGENT, :565-566), game_over [E], done [E,A] as CUDA
        tensors; at E == 1 the leading dimension is dropped."""
        sim = self.sim
        if sim is None:
            raise RuntimeError("call reset() first")
        ext = self._actions_tensor(actions) if self._needs_actions else None
        pol = None
        if self._has_ig and self._mcts is not None:
            # ig_mcts.find_next_action (ig_mcts.py:79-109): team belief update, then Dec-MCTS planning on the host over
            # the GPU rollout evaluator
            sim.explore_team_update()
            done = (sim.state[5].view(torch.int32)[..., 2] & L.F_IS_DONE) != 0
            acts = self._mcts.plan(done.cpu().numpy())
            pol = sim._policy_buffers()
            pol.copy_(torch.from_numpy(acts))
        elif self._has_ig:
            pol = sim.explore_step()
        if self._has_ga3c:
            pol = sim.ga3c_forward(use_tensor_cores=self.ga3c_use_tensor_cores, exact=self.ga3c_exact)
        if self._has_rvo:
            pol = sim.rvo_step(coop=self._rvo_coop, params=self._rvo_params)
        sim.step(ext_actions=ext, policy_actions=pol, dt=dt)
        self.episode_step_number += 1
        self._host_state = None
        # copies: the next step overwrites the simulator's result buffers in place
        rewards, done, game_over = sim.buf["rewards"].clone(), sim.buf["done"].bool(), sim.buf["game_over"].bool()
        if self.config.TRAIN_SINGLE_AGENT:
            rewards = rewards[:, 0]
        if self.E == 1:
            rewards, done, game_over = rewards[0], done[0], game_over[0]
        return BatchedObs(self), rewards, game_over, {"which_agents_done": done}

    def reset_done_envs(self, game_over=None):
        """DummyVecEnv's ``if done: obs = env.reset()`` for the batch: every environment whose ``game_over`` flag is set
        starts a new episode NOW (state, observations, exploration belief, Dec-MCTS team state); with a scenario
        generator the restarted environments draw their next scenario (seed of their own next episode number, like
        _init_agents, collision_avoidance_env.py:403-415), otherwise they restart from their initial condition.
        Returns the mask (bool CUDA tensor [E])."""
        sim = self.sim
        mask = sim.buf["game_over"] if game_over is None else game_over
        mb = mask.bool()
        if self._generator is not None:
            kind, n_ag, ego_pol, ego_dyn, other_dyn, base = self._generator
            self._env_episode += mb.cpu().numpy().astype(np.int64)
            seeds = (base + self._env_episode * self.E + np.arange(self.E) + 1).astype(np.uint32)
            sim.generate_scenario(kind, n_ag, seeds, ego_pol, ego_dyn, other_dyn, out=self._gen_out)
        sim.reset_envs(mask)
        if self._mcts is not None:
            self._mcts.reset_envs(mb.cpu().numpy())
        self._host_state = None
        return mb

    def episode_stats(self):
        """per-step sums over this shard: n_new_collisions, n_new_at_goal, n_new_timeouts, n_game_over_envs,
        reward_sum, n_active_agents, n_agents, 0 (all-reduced across ranks by casim.dist.all_reduce_stats)"""
        return self.sim.buf["stats"]

    def close(self):
        if self.sim is not None:
            self.sim.close()
            self.sim = None


# ---- gym registration (gym_collision_avoidance/__init__.py:6-9) ----
ENV_ID = "CollisionAvoidance-v0"
_REGISTRY = {ENV_ID: BatchedCollisionAvoidanceEnv}


def register():
    """Register ``CollisionAvoidance-v0`` with gym when gym is importable (same id and call as the reference's package
    __init__); ``casim.env.make`` works either way."""
    try:
        from gym.envs.registration import register as _reg      # pragma: no cover - gym is absent from the build image
        _reg(id=ENV_ID, entry_point="casim.env:BatchedCollisionAvoidanceEnv")
        return True
    except Exception:
        return False


def make(env_id=ENV_ID, **kwargs):
    """gym.make(env_id) for this package: ``make('CollisionAvoidance-v0', num_envs=4096)``."""
    return _REGISTRY[env_id](**kwargs)
