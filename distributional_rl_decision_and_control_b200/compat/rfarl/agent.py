from distributional_rl_decision_and_control_b200.agent import *  # noqa: F401,F403
from distributional_rl_decision_and_control_b200.agent import Agent, calculate_huber_loss  # noqa: F401
