"""Instrumentation (C15): device/wall timers, psutil overhead, JSONL metrics."""
from .timers import PhaseTimer, cuda_time_ms  # noqa: F401
from .overhead import OverheadMeter  # noqa: F401
from .log import MetricsLogger  # noqa: F401
from .report import info_passing_table, sweep_table, accuracy_curves, write_csv  # noqa: F401
