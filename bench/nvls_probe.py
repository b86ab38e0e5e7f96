"""Does this box expose NVLS multicast to user kernels?  (torchrun, one process per GPU.)

Uses torch's symmetric-memory allocator only as plumbing: CUDA VMM allocation, handle exchange between the ranks and
`cuMulticastCreate / AddDevice / BindMem` are host-side set-up; what it yields — per-rank peer pointers and ONE multicast
pointer per buffer — is what `fedavg_reduce_nvls_kernel` consumes (`multimem.ld_reduce` / `multimem.st`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from fedb200.parallel import init_distributed

rank, world, local = init_distributed(torch.device("cuda"))
dev = torch.device("cuda", local)
import torch.distributed._symmetric_memory as symm
from torch._C._distributed_c10d import _SymmetricMemory
ok_dev = None
t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
h = symm.rendezvous(t, dist.group.WORLD)
info = {"rank": rank, "world": h.world_size, "multicast_ptr": hex(h.multicast_ptr), "buffer_ptrs": [hex(p) for p in h.buffer_ptrs],
        "signal_pad_size": h.signal_pad_size, "has_multicast_support": ok_dev}
out = [None] * world
dist.all_gather_object(out, info)
if rank == 0:
    for o in out:
        print(o)
    print("NVLS_PROBE", "multicast" if all(int(o["multicast_ptr"], 16) != 0 for o in out) else "no-multicast")
dist.barrier()
dist.destroy_process_group()
