from .simulation_parameters import SimulationParameters  # noqa: F401
from .simulation_runner import SimulationRunner  # noqa: F401
