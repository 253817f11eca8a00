from distributional_rl_decision_and_control_b200.policy.AC_IQN_model import AC_IQN_Policy, Actor, Critic, encoder  # noqa: F401
