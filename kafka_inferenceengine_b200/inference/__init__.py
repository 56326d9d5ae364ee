"""Mirror of the reference's ``kafka.inference`` package
(kafka/inference/__init__.py:1-3): same public names, CUDA behind them."""
from .kf_tools import *      # noqa: F401,F403
from .solvers import *       # noqa: F401,F403
from .utils import *         # noqa: F401,F403
