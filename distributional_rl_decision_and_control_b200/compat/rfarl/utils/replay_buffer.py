from distributional_rl_decision_and_control_b200.utils.replay_buffer import ReplayBuffer  # noqa: F401
