"""``Trainer`` with the reference's constructor and ``learn`` / ``evaluation`` /
``save_evaluation`` / ``save_eval_config`` surface (rfarl/rfarl/policy/trainer.py:7-410) for the
single-environment loop that ``train_RL_agents.py`` drives.  The many-environment loop is
``vector_trainer.VectorTrainer``.

Same schedule as the reference: act -> step -> store -> (every UPDATE_EVERY steps after
learning_starts) train -> (every target_update_interval) target sync -> periodic evaluation over
the frozen scenarios.  One deliberate deviation: when ``rl_agent.training`` is False the
reference `continue`s past its own timestep increment (TRAINER:178-179) and never terminates;
here the counter always advances.
"""
import copy
import json
import os

import numpy as np

_CONTINUOUS = ("AC-IQN", "DDPG", "SAC")


class Trainer():
    def __init__(self,
                 train_env,
                 eval_env,
                 eval_schedule,
                 rl_agent,
                 UPDATE_EVERY=4,
                 learning_starts=2000,
                 target_update_interval=10000,
                 exploration_fraction=0.25,
                 initial_eps=0.6,
                 final_eps=0.05,
                 imitation=False,
                 il_agent=None,
                 batched_evaluation=False):
        self.train_env = train_env
        # True: the frozen scenarios are played as ONE device batch (evaluation.BatchedEvaluator)
        # instead of one after the other; same statistics, no per-step host round trips
        self.batched_evaluation = bool(batched_evaluation)
        self._evaluator = None
        self.eval_env = eval_env
        self.rl_agent = rl_agent
        self.eval_config = []
        self.create_eval_configs(eval_schedule)

        self.UPDATE_EVERY = UPDATE_EVERY
        self.learning_starts = 0 if imitation else learning_starts
        self.target_update_interval = target_update_interval
        self.exploration_fraction = exploration_fraction
        self.initial_eps = initial_eps
        self.final_eps = final_eps
        self.imitation = imitation
        self.il_agent = il_agent
        if self.imitation:
            assert self.il_agent is not None, "Imitation Learning agent not given!"

        self.current_timestep = 0
        self.learning_timestep = 0
        self.eval_timesteps = []
        self.eval_observations = []
        self.eval_actions = []
        self.eval_trajectories = []
        self.eval_rewards = []
        self.eval_successes = []
        self.eval_times = []
        self.eval_energies = []
        self.eval_relations = []

    # ------------------------------------------------------------------ frozen eval scenarios
    def create_eval_configs(self, eval_schedule):
        """TRAINER:63-78: one stored scenario per evaluation episode."""
        self.eval_config.clear()
        for i, num_episode in enumerate(eval_schedule["num_episodes"]):
            for _ in range(num_episode):
                self.eval_env.num_robots = eval_schedule["num_robots"][i]
                self.eval_env.num_cores = eval_schedule["num_cores"][i]
                self.eval_env.num_obs = eval_schedule["num_obstacles"][i]
                self.eval_env.min_start_goal_dis = eval_schedule["min_start_goal_dis"][i]
                self.eval_env.reset()
                self.eval_config.append(self.eval_env.episode_data())

    def save_eval_config(self, directory):
        with open(os.path.join(directory, "eval_configs.json"), "w+") as f:
            json.dump(self.eval_config, f)

    # ------------------------------------------------------------------ acting helper
    def _act(self, state, eps, use_eval):
        agent = self.rl_agent
        kind = agent.agent_type
        if kind == "AC-IQN":
            return agent.act_ac_iqn(state, eps, use_eval=use_eval)
        if kind == "IQN":
            return agent.act_iqn(state, eps, use_eval=use_eval)[0]
        if kind == "DDPG":
            return agent.act_ddpg(state, eps, use_eval=use_eval)
        if kind == "DQN":
            return agent.act_dqn(state, eps, use_eval=use_eval)
        if kind == "SAC":
            return agent.act_sac(state, eps, use_eval=use_eval)
        if kind == "Rainbow":
            return agent.act_rainbow(state, eps, use_eval=use_eval)
        raise RuntimeError("Agent type not implemented!")

    # ------------------------------------------------------------------ training loop
    def learn(self, total_timesteps, eval_freq, eval_log_path, verbose=True):
        env, agent = self.train_env, self.rl_agent
        is_continuous_action = agent.agent_type in _CONTINUOUS
        states, _, _ = env.reset()
        ep_rewards = np.zeros(len(env.robots))
        ep_deactivated_t = [-1] * len(env.robots)
        ep_length = 0
        ep_num = 0
        eps = self.final_eps

        while self.current_timestep <= total_timesteps:
            if not self.imitation:
                eps = self.linear_eps(total_timesteps)

            actions = []
            for i, rob in enumerate(env.robots):
                if rob.deactivated:
                    actions.append(None)
                elif self.imitation:
                    actions.append(self.il_agent.act(states[i]))
                else:
                    actions.append(self._act(states[i], eps, use_eval=False))

            next_states, rewards, dones, infos = env.step(actions, is_continuous_action)

            for i, rob in enumerate(env.robots):  # TRAINER:157-170
                if rob.deactivated:
                    continue
                ep_rewards[i] += agent.GAMMA ** ep_length * rewards[i]
                if agent.training:
                    if agent.agent_type == "Rainbow":  # TRAINER:163-164
                        agent.memory.append(states[i], actions[i], rewards[i], dones[i])
                    else:
                        agent.memory.add((states[i], actions[i], rewards[i], next_states[i], dones[i]))
                if rob.collision or rob.reach_goal:
                    rob.deactivated = True
                    ep_deactivated_t[i] = ep_length

            end_episode = (ep_length >= 1000) or env.check_all_deactivated()

            if self.current_timestep >= self.learning_starts and agent.training:
                if self.current_timestep % self.UPDATE_EVERY == 0:
                    num_elements = agent.memory.transitions.num_elements() if agent.agent_type == "Rainbow" \
                        else agent.memory.size()  # TRAINER:184-185
                    if num_elements > agent.BATCH_SIZE:
                        agent.train()
                if self.current_timestep % self.target_update_interval == 0:
                    agent.soft_update()
                if self.current_timestep == self.learning_starts or self.current_timestep % eval_freq == 0:
                    self.evaluation()
                    self.save_evaluation(eval_log_path)
                    agent.save_latest_model(eval_log_path)

            if end_episode:
                ep_num += 1
                if verbose:
                    self._print_episode(ep_length, ep_num, eps, total_timesteps, infos, ep_rewards,
                                        ep_deactivated_t)
                states, _, _ = env.reset()
                ep_rewards = np.zeros(len(env.robots))
                ep_deactivated_t = [-1] * len(env.robots)
                ep_length = 0
            else:
                states = next_states
                ep_length += 1
            self.current_timestep += 1

    def _print_episode(self, ep_length, ep_num, eps, total_timesteps, infos, ep_rewards, ep_deactivated_t):
        print("======== IL Episode Info ========" if self.imitation else "======== RL Episode Info ========")
        print("current ep_length: ", ep_length)
        print("current ep_num: ", ep_num)
        if not self.imitation:
            print("current exploration rate: ", eps)
        print("current timesteps: ", self.current_timestep)
        print("total timesteps: ", total_timesteps)
        print("======== Episode Info ========\n")
        print("======== Robots Info ========")
        for i in range(len(infos)):
            info = infos[i]["state"]
            if info.startswith("deactivated after"):
                print(f"Robot {i} ep reward: {ep_rewards[i]:.2f}, {info} at step {ep_deactivated_t[i]}")
            else:
                print(f"Robot {i} ep reward: {ep_rewards[i]:.2f}, {info}")
        print("======== Robots Info ========\n")

    def linear_eps(self, total_timesteps):
        """TRAINER:257-264."""
        progress = self.current_timestep / total_timesteps
        if progress < self.exploration_fraction:
            r = progress / self.exploration_fraction
            return self.initial_eps + r * (self.final_eps - self.initial_eps)
        return self.final_eps

    # ------------------------------------------------------------------ evaluation (TRAINER:266-392)
    def evaluation(self):
        """Greedy roll-outs over the frozen scenarios; an episode ends after 1000 steps, on ANY
        collision, or when every robot is deactivated (TRAINER:344)."""
        env, agent = self.eval_env, self.rl_agent
        is_continuous_action = agent.agent_type in _CONTINUOUS
        data = {k: [] for k in ("observations", "actions", "trajectories", "rewards", "successes",
                                "times", "energies", "relations")}
        if self.batched_evaluation:
            from ..evaluation import BatchedEvaluator
            if self._evaluator is None:
                self._evaluator = BatchedEvaluator(self.eval_config, device=env.device, seed=env.seed)
            res = self._evaluator.run(agent)
            for k in ("rewards", "successes", "times", "energies", "actions", "trajectories"):
                data[k] = res[k]
            data["observations"] = [[[] for _ in t] for t in res["trajectories"]]  # not recorded in batch mode
            data["relations"] = [[[] for _ in t] for t in res["trajectories"]]
        for idx, config in enumerate([] if self.batched_evaluation else self.eval_config):
            print(f"Evaluating episode {idx}")
            state, _, _ = env.reset_with_eval_config(config)
            n = len(env.robots)
            rewards, times, energies = [0.0] * n, [0.0] * n, [0.0] * n
            length = 0
            end_episode = False
            while not end_episode:
                action = [None if rob.deactivated else self._act(state[i], 0.0, use_eval=True)
                          for i, rob in enumerate(env.robots)]
                state, reward, done, info = env.step(action, is_continuous_action)
                for i, rob in enumerate(env.robots):
                    if rob.deactivated:
                        continue
                    rewards[i] += agent.GAMMA ** length * reward[i]
                    times[i] += rob.dt * rob.N
                    energies[i] += rob.compute_step_energy_cost()
                    if rob.collision or rob.reach_goal:
                        rob.deactivated = True
                end_episode = (length >= 1000) or env.check_any_collision() or env.check_all_deactivated()
                length += 1
            data["observations"].append([copy.deepcopy(rob.observation_history) for rob in env.robots])
            data["actions"].append([copy.deepcopy(rob.action_history) for rob in env.robots])
            data["trajectories"].append([copy.deepcopy(rob.trajectory) for rob in env.robots])
            data["rewards"].append(np.mean(rewards))
            data["successes"].append(True if env.check_all_reach_goal() else False)
            data["times"].append(np.mean(times))
            data["energies"].append(np.mean(energies))
            data["relations"].append([[] for _ in range(n)])

        avg_r = np.mean(data["rewards"])
        success_rate = np.sum(data["successes"]) / len(data["successes"])
        won = np.where(np.array(data["successes"]) == 1)[0]
        avg_t = None if won.shape[0] == 0 else np.mean(np.array(data["times"])[won])
        avg_e = None if won.shape[0] == 0 else np.mean(np.array(data["energies"])[won])
        print("++++++++ Evaluation Info ++++++++")
        print(f"Avg cumulative reward: {avg_r:.2f}")
        print(f"Success rate: {success_rate:.2f}")
        if avg_t is not None:
            print(f"Avg time: {avg_t:.2f}")
            print(f"Avg energy: {avg_e:.2f}")
        print("++++++++ Evaluation Info ++++++++\n")

        self.eval_timesteps.append(self.current_timestep)
        self.eval_observations.append(data["observations"])
        self.eval_actions.append(data["actions"])
        self.eval_trajectories.append(data["trajectories"])
        self.eval_rewards.append(data["rewards"])
        self.eval_successes.append(data["successes"])
        self.eval_times.append(data["times"])
        self.eval_energies.append(data["energies"])
        self.eval_relations.append(data["relations"])

    def save_evaluation(self, eval_log_path):
        """evaluations.npz with the reference's keys (TRAINER:394-410)."""
        obj = lambda v: np.array(v, dtype=object)
        np.savez(os.path.join(eval_log_path, "evaluations.npz"),
                 timesteps=obj(self.eval_timesteps), observations=obj(self.eval_observations),
                 actions=obj(self.eval_actions), trajectories=obj(self.eval_trajectories),
                 rewards=obj(self.eval_rewards), successes=obj(self.eval_successes),
                 times=obj(self.eval_times), energies=obj(self.eval_energies),
                 relations=obj(self.eval_relations))
