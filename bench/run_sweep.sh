#!/bin/bash
# Client-count sweep of the reference (for NUM_CLIENTS in [5, 10, 20], 20 rounds; serverless_cancer_biobert_allclients.py:40-41)
# for server and serverless mode on both dataset shapes, on ONE GPU (peers time-share it, as in the reference's single
# process).  Produces gpurun_out/sweep/<mode>_<dataset>.json = the latency / accuracy / memory-vs-workers table and the
# accuracy-vs-round curves that BASELINE.md sets beside the notebook literals.   Usage: bench/run_sweep.sh [rounds] [model] [lr]
R=${1:-20}; M=${2:-biobert}; LR=${3:-5e-5}
OUT=gpurun_out/sweep; mkdir -p $OUT
for ds in imdb medical_transcriptions; do
  for mode in server serverless; do
    python ${mode}_IID.py --model $M --dataset $ds --rounds $R --sweep 5,10,20 --lr $LR --ledger --metrics $OUT/${mode}_${ds}.jsonl --quiet \
        > $OUT/${mode}_${ds}.log 2>&1
    grep '^{' $OUT/${mode}_${ds}.log | tail -1 > $OUT/${mode}_${ds}.json
    echo "$mode $ds: $(cut -c1-300 $OUT/${mode}_${ds}.json)"
  done
done
