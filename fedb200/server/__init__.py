"""Server mode: FedAvg strategy (C8), metric aggregation (C9) and the simulation driver (C10)."""
from .fedavg import FedAvg, weighted_average, aggregate_weighted  # noqa: F401
from .simulation import start_simulation, ServerConfig, History  # noqa: F401
