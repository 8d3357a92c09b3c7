"""Drop-in for the reference's actor.py (main.py:6-8 imports it by this name): re-exports the
B200 classes.  Put this directory before the reference on PYTHONPATH."""
from apex_b200.actor import *            # noqa: F401,F403
from apex_b200.actor import N_Step_Transition   # noqa: F401
from apex_b200.actor import Actor, ExperienceBuffer, Transition   # noqa: F401,E402
