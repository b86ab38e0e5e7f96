"""Launch every kernel family of libfedb200.so once on ONE GPU, for `ncu --set full` (profiles/ncu/<kernel>.txt).

The multi-GPU kernels (fedavg_reduce, gossip_mix, gossip_mix_async, fused Gram, exchange) are launched in their one-rank
form: same code, the "peer" pointers are this GPU's own buffers and the cross-rank barriers involve only the rank's own flag
words, so ncu (which replays each kernel ~40 times and cannot run under torchrun) can profile them.  NVLink counters are not
exposed on this box (the six `ctc__*` link metrics come back "n/a"); link traffic is reported from CUDA-event times and the
algorithmic byte counts in profiles/comm_*.log instead.

    python bench/ncu_kernels.py            # plain run (must exit 0 before the ncu run)
    ncu --set full --clock-control none --import-source on -o gpurun_out/prof_all python bench/ncu_kernels.py
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from fedb200.config import FedConfig, ModelConfig
from fedb200.data import make_synthetic
from fedb200.data.dataset import collate_packed
from fedb200.models.fused_bert import FusedRunner
from fedb200.ops import _lib, cuda_ops

dev = torch.device("cuda")
D = 108_312_064 if os.environ.get("FEDB200_NCU_SMALL", "0") != "1" else 8_388_608
stream = lambda: torch.cuda.current_stream().cuda_stream
arr = lambda xs: C.cast((C.c_void_p * len(xs))(*[int(x) for x in xs]), C.c_void_p)

# ---- local trainer: one training step + one evaluation step (K8-K13: GEMMs, attention, LayerNorm, embed, head, AdamW)
fed = FedConfig()
fed.model = ModelConfig.preset("biobert", num_labels=2)
run = FusedRunner(fed.model, fed, "cuda", seed=0)
ds = make_synthetic("imdb", "train", fed.model.vocab_size, 512, rows=64)
batch = collate_packed(ds, np.arange(32), longest_first=True)
batch = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
run.train_step(batch)
run.eval_step(batch)
torch.cuda.synchronize()
del run

# ---- federated kernels, one-rank forms ---------------------------------------------------------------------
x = torch.randn(D, device=dev)
y = torch.empty(D, device=dev)
shadow = torch.empty(D, dtype=torch.bfloat16, device=dev)
flags = torch.zeros(128, dtype=torch.int32, device=dev)
# K1 fused FedAvg reduce (two-shot form, world = 1) + one-shot small form + slot exchange
_lib.call("fedb200_fedavg_reduce_ex", arr([x.data_ptr()]), arr([y.data_ptr()]), arr([shadow.data_ptr()]), arr([flags.data_ptr()]),
          0, 1, D, 1.0, 0.0, 1, 2000, 0, 0, stream())
sm_a, sm_o = torch.randn(4096, device=dev), torch.empty(4096, device=dev)
_lib.call("fedb200_fedavg_reduce_ex", arr([sm_a.data_ptr()]), arr([sm_o.data_ptr()]), None, arr([flags.data_ptr()]),
          0, 1, 4096, 1.0, 0.0, 1, 2000, 1, 0, stream())
tab = torch.zeros(1, 192, dtype=torch.int32, device=dev)
_lib.call("fedb200_peer_exchange", arr([tab.data_ptr()]), arr([flags.data_ptr()]), 0, 1, 192 * 4, 1, 2000, stream())
# K2 fused gossip mix: 3 sources -> 2 rows
srcs = [x, y, torch.randn(D, device=dev)]
outs = [torch.empty(D, device=dev) for _ in range(2)]
w = np.ascontiguousarray(np.array([[0.5, 0.25, 0.25], [0.2, 0.4, 0.4]], dtype=np.float32))
_lib.call("fedb200_gossip_mix", arr([s.data_ptr() for s in srcs]), w.ctypes.data_as(C.c_void_p), 3, arr([o.data_ptr() for o in outs]), 2,
          arr([0]), arr([0]), 0, 0, arr([flags.data_ptr()]), 0, 1, D, 1, 2000, stream())
# K3 asynchronous mixer: publish two versions with landing flags, then mix from the observed versions
R, chunk = 3, 1 << 20
nch = (D + chunk - 1) // chunk
ring = torch.empty(R, D, device=dev)
cfl = torch.zeros(R, nch, dtype=torch.int32, device=dev)
pub = torch.zeros(4, dtype=torch.int32, device=dev)
for e in (1, 2):
    pub[0] = e
    cuda_ops.copy_flagged(x, ring[e % R], chunk, cfl[e % R], e)
decision = torch.zeros(16, dtype=torch.int32, device=dev)
log = torch.zeros(8, dtype=torch.int32, device=dev)
err = torch.zeros(4, dtype=torch.int32, device=dev)
from fedb200.p2p.observed import stale_table
cuda_ops.gossip_mix_async(y, outs[0], [ring, ring], [cfl, cfl], [pub, pub], [1, 2], None, decision, log, err.data_ptr(), R, chunk, 3, 1,
                          stale_table(0.5, 0.5))
# K4 fused Gram (8 clients, segment-stacked) + K5 PageRank
N = 8
lay = cuda_ops.gram_layout(D, 1, N)
deltas = torch.zeros(N, lay["D_pad"], dtype=torch.bfloat16, device=dev)
deltas[:, :D] = (torch.randn(N, D, device=dev) * 0.01).bfloat16()
part, res = torch.zeros(2 * N * N, device=dev), torch.empty(N, N, device=dev)
_lib.call("fedb200_gram_fused_seg", arr([deltas.data_ptr()]), 1, N, lay["seg_len"], lay["nseg_total"], 0, lay["S"],
          arr([part.data_ptr()]), arr([flags.data_ptr()]), res.data_ptr(), 0, 1, 2000, 0, stream())
cuda_ops.pagerank(res.contiguous(), mode=1, low_only=True)
del deltas
# K6 SHA-256 fused hash + copy, Merkle reduction, K7 block validation
leaves = torch.empty((D * 4 + 4095) // 4096, 32, dtype=torch.uint8, device=dev)
cuda_ops.sha256_leaves(x, y, 4096, leaves)
scratch = torch.empty(leaves.shape[0] // 2 + 2, 32, dtype=torch.uint8, device=dev)
root = torch.empty(32, dtype=torch.uint8, device=dev)
cuda_ops.merkle_root(leaves, scratch, root)
from fedb200.ledger import Ledger
led = Ledger(chunk_bytes=4096)
for i in range(64):
    led.record_update(i // 8, i % 8, 0xFFFFFFFF, np.full(16, i, dtype=np.float32).tobytes(), 1, [i % 8])
hdr = torch.from_numpy(led.headers_array()).to(dev)
hashes = torch.from_numpy(np.frombuffer(b"".join(led.hashes), dtype=np.uint8).copy()).to(dev)
hashes = hashes.view(64, 32)
bad_idx, _ = cuda_ops.block_validate(hdr, hashes)
# flat-arena elementwise kernels
cuda_ops.mix_n([x, y], [0.5, 0.5], outs[0], shadow)
cuda_ops.cast_bf16(x, shadow)
cuda_ops.delta_bf16(x, y, shadow)
torch.cuda.synchronize()
assert int(flags[18]) == 0 and int(err[0]) == 0 and bad_idx == -1 and int(log[2]) == 0, (flags[18], err, bad_idx, log)
print("ncu_kernels: ok")
