from distributional_rl_decision_and_control_b200.policy.DQN_model import DQN_Policy, encoder  # noqa: F401
