"""Anomalous-node detectors (C18-C21), the cosine/PageRank update scorer (K4/K5 host oracle)
and the fault injector."""
from .pagerank import pagerank, pagerank_anomalies  # noqa: F401
from .modz import modified_z_scores, modz_anomalies  # noqa: F401
from .dbscan import dbscan_1d, dbscan_anomalies  # noqa: F401
from .community import greedy_modularity_communities, community_anomalies  # noqa: F401
from .cosine import cosine_gram, gram_to_edge_weights, similarity_to_transition, score_updates  # noqa: F401
from .inject import inject_anomaly, AnomalyInjector  # noqa: F401
from .detect import detect  # noqa: F401
