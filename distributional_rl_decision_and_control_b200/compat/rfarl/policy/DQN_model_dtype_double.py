from distributional_rl_decision_and_control_b200.policy.DQN_model_dtype_double import DQN_Policy  # noqa: F401
