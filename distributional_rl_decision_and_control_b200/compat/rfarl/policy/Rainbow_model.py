from distributional_rl_decision_and_control_b200.policy.Rainbow_model import Rainbow_Policy, NoisyLinear, encoder  # noqa: F401
