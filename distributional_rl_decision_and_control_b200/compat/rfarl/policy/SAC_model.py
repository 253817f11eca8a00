from distributional_rl_decision_and_control_b200.policy.SAC_model import SAC_Policy, Actor, Critic, encoder  # noqa: F401
