from distributional_rl_decision_and_control_b200.utils.replay_memory_rainbow import ReplayMemory  # noqa: F401
