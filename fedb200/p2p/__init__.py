"""Peer-to-peer layer: network graph model (C17), information-passing-time model (C22),
path optimiser (C23), gossip mixing (C11/C12, K2) and the asynchronous engine (C25, K3)."""
from .topology import Topology, REFERENCE_BANDWIDTH_MBPS  # noqa: F401
from .path_opt import (shortest_paths, info_passing_time, optimise_route, RoutePlan)  # noqa: F401
from .gossip import mixing_matrix, apply_mask, mix_flat, RelayPlan  # noqa: F401
from .async_engine import AsyncGossipEngine, staleness_weight  # noqa: F401
from .schedule import build_async_schedule, model_durations, AsyncSchedule, ScheduledEvent  # noqa: F401
