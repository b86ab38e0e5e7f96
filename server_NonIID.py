#!/usr/bin/env python
"""server_NonIID — FedAvg with non-IID client splits: the reference's contiguous shards
(src/Servercase/server_NonIID_IMDB.py:65-101) or ``--split dirichlet`` label skew."""
from fedb200.entry import main

if __name__ == "__main__":
    main("server", "shard", __doc__)
