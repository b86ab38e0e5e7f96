#!/bin/bash
# The named configurations of BASELINE.json.  Usage: bench/run_configs.sh [N] [model-for-2..4] [model-for-5]
#   c1 server_IID, BERT-base, 4 clients, CPU plumbing (no GPU; runs anywhere)          ONLY="c1"
#   c2 serverless_IID, 8 peers, synchronous gossip over NVSwitch
#   c3 serverless_NonIID async, BioBERT medical-transcription shape, 2 poisoned peers (sign-flipped updates), on-line
#      cosine / PageRank elimination with the mask kept on the device
#   c4 server_NonIID, 32 clients (Dirichlet), ledgered FedAvg, token-balanced placement
#   c5 serverless async scale-up: BERT-large, 64 peers, observed-version gossip with bounded fan-in, executed routes
N=${1:-8}; M=${2:-biobert}; ML=${3:-bert-large}
OUT=gpurun_out/configs; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1"
ONLY=${ONLY:-c2 c3 c4 c5}
R2=${ROUNDS:-3}
run() { name=$1; shift; case " $ONLY " in *" ${name%%_*} "*) ;; *) return;; esac; echo "=== $name: $*"; ( time timeout 1500 "$@" ) > $OUT/$name.log 2>&1; grep '^{"global_accuracies\|^{"' $OUT/$name.log | tail -1 > $OUT/$name.json; grep -E "^real" $OUT/$name.log; cut -c1-400 $OUT/$name.json; }
run c1_server_IID_cpu python server_IID.py --model bert-base --clients 4 --rounds 1 --device cpu --backend torch --train-per-client 32 --test-per-client 32 \
    --max-seq-len 128 --ledger --metrics $OUT/c1.jsonl --quiet
run c2_serverless_IID_sync $TR --master-port 29551 serverless_IID.py --model $M --clients $N --rounds $R2 --ledger --metrics $OUT/c2.jsonl --quiet
run c3_serverless_NonIID_async_poisoned $TR --master-port 29552 serverless_NonIID.py --model $M --dataset medical_transcriptions --clients $N --rounds 4 --async \
    --topology full --inject-anomalies 2 --inject-kind signflip --anomaly-detector cosine_pagerank --async-score-every 1 --ledger --metrics $OUT/c3.jsonl --quiet
run c4_server_NonIID_32clients_ledger $TR --master-port 29553 server_NonIID.py --model $M --clients $((4*N)) --rounds 2 --split dirichlet --ledger --metrics $OUT/c4.jsonl --quiet
run c5_serverless_async_large_64peers $TR --master-port 29554 serverless_IID.py --model $ML --clients $((8*N)) --rounds 2 --async --topology full --async-fanin 4 --ledger --metrics $OUT/c5.jsonl --quiet
