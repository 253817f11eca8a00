from distributional_rl_decision_and_control_b200.policy.AC_IQN_model_dtype_double import Actor  # noqa: F401
