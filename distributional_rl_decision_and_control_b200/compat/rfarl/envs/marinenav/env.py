from distributional_rl_decision_and_control_b200.envs.marinenav.env import Core, Obstacle, MarineNavEnv3  # noqa: F401
