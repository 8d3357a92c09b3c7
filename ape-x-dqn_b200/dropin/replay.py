"""Drop-in for the reference's replay.py (main.py:6-8 imports it by this name): re-exports the
B200 classes.  Put this directory before the reference on PYTHONPATH."""
from apex_b200.replay import *            # noqa: F401,F403
from apex_b200.replay import N_Step_Transition   # noqa: F401
from apex_b200.replay import ReplayMemory   # noqa: F401,E402
