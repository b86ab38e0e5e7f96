#!/bin/bash
# The named configurations of BASELINE.json on N GPUs of one box (default 8).  Usage: bench/run_configs.sh [N] [model-for-2..4] [model-for-5]
N=${1:-8}; M=${2:-biobert}; ML=${3:-bert-large}
OUT=gpurun_out/configs; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
ONLY=${ONLY:-c2 c3 c4 c5}
run() { name=$1; shift; case " $ONLY " in *" ${name%%_*} "*) ;; *) return;; esac; echo "=== $name: $*"; ( time timeout 900 "$@" ) > $OUT/$name.log 2>&1; grep '^{"global_accuracies\|^{"' $OUT/$name.log | tail -1 > $OUT/$name.json; grep -E "^real" $OUT/$name.log; cut -c1-400 $OUT/$name.json; }
run c2_serverless_IID_sync $TR --master-port 29551 serverless_IID.py --model $M --clients $N --rounds 3 --ledger --metrics $OUT/c2.jsonl --quiet
run c3_serverless_NonIID_async_pagerank $TR --master-port 29552 serverless_NonIID.py --model $M --dataset medical_transcriptions --clients $N --rounds 3 --async \
    --topology reference10 --inject-anomalies 2 --inject-kind slowlink --anomaly-detector pagerank --ledger --async-epoch-s 20 --metrics $OUT/c3.jsonl --quiet
run c4_server_NonIID_32clients_ledger $TR --master-port 29553 server_NonIID.py --model $M --clients $((4*N)) --rounds 2 --split dirichlet --ledger --metrics $OUT/c4.jsonl --quiet
run c5_serverless_async_large_64peers $TR --master-port 29554 serverless_IID.py --model $ML --clients $((8*N)) --rounds 2 --async --topology full --ledger --async-min-latency-s 0.05 --metrics $OUT/c5.jsonl --quiet
