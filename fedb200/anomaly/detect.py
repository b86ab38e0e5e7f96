"""Dispatch over the four graph detectors of the notebooks (C18-C21) on a ``Topology``."""
from __future__ import annotations

from typing import List

from ..p2p.topology import Topology
from .community import community_anomalies
from .dbscan import dbscan_anomalies
from .modz import modz_anomalies
from .pagerank import pagerank_anomalies


def detect(topo: Topology, method: str) -> List[int]:
    if method in ("none", "", None):
        return []
    if method == "pagerank":
        return pagerank_anomalies(topo.edge_weights())[0]
    if method == "modz":
        return modz_anomalies(topo.weighted_degree_undirected())[0]
    if method == "dbscan":
        return dbscan_anomalies(topo.weighted_degree_undirected())[0]
    if method in ("community", "louvain"):
        return community_anomalies(topo.edge_weights())
    raise ValueError(f"unknown detector {method!r}")
