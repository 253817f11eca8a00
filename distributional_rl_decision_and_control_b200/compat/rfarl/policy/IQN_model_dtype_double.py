from distributional_rl_decision_and_control_b200.policy.IQN_model_dtype_double import IQN_Policy  # noqa: F401
