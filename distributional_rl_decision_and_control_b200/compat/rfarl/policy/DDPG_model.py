from distributional_rl_decision_and_control_b200.policy.DDPG_model import DDPG_Policy, Actor, Critic, encoder  # noqa: F401
