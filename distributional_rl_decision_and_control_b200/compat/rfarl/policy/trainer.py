from distributional_rl_decision_and_control_b200.policy.trainer import Trainer  # noqa: F401
