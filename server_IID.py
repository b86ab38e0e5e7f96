#!/usr/bin/env python
"""server_IID — FedAvg with IID client subsets (reference: src/Servercase/server_IID_IMDB.py,
src/Servercase/server_iid_medical_transcirptions.py).  ``torchrun --nproc-per-node 8 server_IID.py``
spreads the clients over the GPUs of one box."""
from fedb200.entry import main

if __name__ == "__main__":
    main("server", "iid", __doc__)
