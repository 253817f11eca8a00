from distributional_rl_decision_and_control_b200.envs.marinenav.vehicles.wamv import Perception, Robot  # noqa: F401
