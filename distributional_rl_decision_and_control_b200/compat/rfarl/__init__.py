"""Import shim: puts this repo's host layer under the reference's package name ``rfarl`` so
``rfarl/rfarl/scripts/train_RL_agents.py`` and ``virelex/.../run_vrx_experiments.py`` run
unchanged with this directory's parent (``.../compat``) first on PYTHONPATH."""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)
