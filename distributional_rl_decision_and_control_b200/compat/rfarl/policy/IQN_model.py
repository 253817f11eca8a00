from distributional_rl_decision_and_control_b200.policy.IQN_model import IQN_Policy, encoder  # noqa: F401
