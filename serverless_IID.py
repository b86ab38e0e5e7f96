#!/usr/bin/env python
"""serverless_IID — peer-to-peer gossip averaging with IID resampling every round (reference:
src/Serverlesscase/serverless_IID_IMDB.py, Serverless_iid_Medical_transcriptions.py,
serverless_caner_classification_iid.py, serverless_covid_iid.py; ``--sweep 5,10,20`` is
serverless_cancer_biobert_allclients.py)."""
from fedb200.entry import main

if __name__ == "__main__":
    main("serverless", "iid", __doc__)
