"""Flattened-observation wrappers and the vectorised-env adapter.

Mirrors gym_collision_avoidance/envs/wrappers.py (``MultiagentFlattenDictWrapper``, ``FlattenDictWrapper``,
``MultiagentDummyVecEnv``) and experiments/src/env_utils.py:7-39 (``create_env``): same class and method names, same
index bookkeeping (``observation_indices``), but ONE batched environment on the GPU takes the place of the list of
Python envs, and the flattening is one kernel (casim_flatten_obs) producing the whole ``[num_envs, size]`` array.
"""
import numpy as np
import torch

from . import layout as L


class MultiagentFlattenDictWrapper:
    """envs/wrappers.py:8-95.  ``observation()`` returns a float32 CUDA tensor [num_envs, size] (a 1-D [size] array at
    num_envs == 1, like the reference); the index helpers are the reference's."""

    def __init__(self, env, dict_keys, max_num_agents=None):
        self.env = env
        self.dict_keys = list(dict_keys)
        self.max_num_agents = int(max_num_agents if max_num_agents is not None else env.A)
        if self.max_num_agents > env.A:
            raise ValueError("max_num_agents %d exceeds the env's agent slots %d" % (self.max_num_agents, env.A))
        self._widths = None
        self.observation_indices = None

    # the widths come from the library once the sim exists (sensor rows / beams / submap are config-dependent)
    def _build(self):
        sim = self.env.sim
        if sim is None:
            raise RuntimeError("reset() the environment first")
        self._widths = [sim.flat_row_size([k]) for k in self.dict_keys]
        self.observation_indices = {}
        size = 0
        for agent in range(self.max_num_agents):        # wrappers.py:19-29
            self.observation_indices[agent] = {}
            lo = size
            for key, w in zip(self.dict_keys, self._widths):
                self.observation_indices[agent][key] = [size, size + w]
                size += w
            self.observation_indices[agent]['BOUNDS'] = [lo, size]
        self.size = size
        self.single_agent_size = size // max(self.max_num_agents, 1)
        from . import spaces as _spaces                                  # wrappers.py:31-35
        self.observation_space = _spaces.Box(-np.inf, np.inf, shape=(size,), dtype='float32')
        self.single_agent_observation_space = _spaces.Box(-np.inf, np.inf, shape=(self.single_agent_size,), dtype='float32')
        self.dict_observation_space = self.env.observation_space
        self.action_space = self.env.action_space

    def reset(self):
        self.env.reset()
        self._build()
        return self.observation()

    def step(self, actions=None, dt=None):
        _, rewards, game_over, info = self.env.step(actions, dt)
        return self.observation(), rewards, game_over, info

    def observation(self, observation=None):
        """wrappers.py:37-45 for every environment of the batch."""
        if self.observation_indices is None:
            self._build()
        sim = self.env.sim
        flat = sim.flatten_obs(self.dict_keys)                                   # [E, A*S]
        if self.max_num_agents < sim.A:
            flat = flat.view(sim.E, sim.A, -1)[:, :self.max_num_agents].reshape(sim.E, -1)
        return flat[0] if sim.E == 1 else flat

    def observationArrayToDict(self, observation_array):  # noqa: N802 (reference name, wrappers.py:47-55)
        obs = {}
        for agent in range(self.max_num_agents):
            obs[agent] = {}
            for key in self.dict_keys:
                a, b = self.observation_indices[agent][key]
                obs[agent][key] = observation_array[a:b]
        return obs

    def multiEnvObservationArrayToDict(self, observation_array):  # noqa: N802 (reference name, wrappers.py:58-73)
        """[num_envs, size] array -> object array [num_envs, max_num_agents] of per-agent dicts, filled only for the
        agents whose 'use_ppo' entry is set, values reshaped to the observation-space shapes -- the reference's
        semantics; index arithmetic on a host copy (this is the consumer side, not the step path)."""
        a = observation_array.detach().cpu().numpy() if isinstance(observation_array, torch.Tensor) else np.asarray(observation_array)
        num_envs = a.shape[0]
        dict_obs = np.empty((num_envs, self.max_num_agents), dtype=dict)
        spaces = self.env.observation_space.spaces
        for env in range(num_envs):
            for agent in range(self.max_num_agents):
                lo, hi = self.observation_indices[agent]['use_ppo']
                if not a[env, lo:hi]:
                    continue
                dict_obs[env][agent] = {}
                for key in self.dict_keys:
                    lo, hi = self.observation_indices[agent][key]
                    v = a[env, lo:hi]
                    shp = spaces[key].shape if key in spaces and int(np.prod(spaces[key].shape)) == v.size else v.shape
                    dict_obs[env][agent][key] = v.reshape(shp)
        return dict_obs

    def singleAgentObservationArrayToDict(self, observation_array, agent):  # noqa: N802 (wrappers.py:75-84)
        out = []
        for env in range(observation_array.shape[0]):
            out.append({key: observation_array[env, self.observation_indices[agent][key][0]:
                                               self.observation_indices[agent][key][1]] for key in self.dict_keys})
        return out

    def keyToArrayInds(self, key):  # noqa: N802 (wrappers.py:86-90)
        return [self.observation_indices[agent][key] for agent in range(self.max_num_agents)]

    def singleAgentObservationArray(self, observation_array, agent):  # noqa: N802 (wrappers.py:92-93)
        a, b = self.observation_indices[agent]['BOUNDS']
        return observation_array[..., a:b]

    def singleAgentObservationInds(self, agent):  # noqa: N802 (wrappers.py:95-96)
        return self.observation_indices[agent]['BOUNDS']


class FlattenDictWrapper(MultiagentFlattenDictWrapper):
    """envs/wrappers.py:97-99: agent 0 only."""

    def __init__(self, env, dict_keys):
        super().__init__(env, dict_keys, max_num_agents=1)


class BatchedVecEnv:
    """What MultiagentDummyVecEnv([make_env] * n_envs) (envs/wrappers.py:101-106, env_utils.py:29-33) gives a trainer:
    ``reset() -> obs [num_envs, size]``, ``step(actions) -> (obs, rews, dones, infos)`` with DummyVecEnv semantics
    (stable_baselines dummy_vec_env.step_wait): an environment that reports ``done`` is reset INSIDE the same call, the
    returned observation row is the new episode's first observation, and the finished episode's last observation is
    handed over as ``infos['terminal_observation']`` (rows of the environments in ``infos['terminal_mask']``).
    The restart is one masked launch (casim_reset_envs) -- no host loop over environments; with a scenario generator
    the restarted environments draw their next scenario, otherwise they restart from their initial condition."""

    def __init__(self, wrapped, keep_terminal_obs=True):
        self.wrapped = wrapped
        self.env = wrapped.env
        self.num_envs = self.env.E
        self.keep_terminal_obs = keep_terminal_obs
        self.env.cfg.auto_reset = 0          # restarts are explicit here (after the step), not folded into the next step

    def reset(self):
        if self.env.sim is not None:
            self.env.sim.set_config(self.env.cfg)
        return self.wrapped.reset()

    def step(self, actions=None):
        obs, rewards, game_over, info = self.wrapped.step(actions)
        info = dict(info)
        if bool(game_over.any()):
            if self.keep_terminal_obs:
                info["terminal_observation"] = obs.clone()
            info["terminal_mask"] = game_over.reshape(-1).clone()
            self.env.reset_done_envs()
            obs = self.wrapped.observation()
        return obs, rewards, game_over, info

    def step_async(self, actions):
        self._pending = actions

    def step_wait(self):
        return self.step(self._pending)

    def close(self):
        self.env.close()
