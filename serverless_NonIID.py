#!/usr/bin/env python
"""serverless_NonIID — peer-to-peer gossip with contiguous-shard (or Dirichlet) client data
(reference: src/Serverlesscase/serverless_NonIID_IMDB.py,
Serverless_NonIID_Medical_transcriptions.py).  ``--async`` enables staleness-weighted
asynchronous mixing; ``--anomaly-detector cosine_pagerank --inject-anomalies 2`` is the
BASELINE.json config 3."""
from fedb200.entry import main

if __name__ == "__main__":
    main("serverless", "shard", __doc__)
