"""Mirror of kafka/observation_operators/__init__.py."""
from .sar_forward_model import *      # noqa: F401,F403
