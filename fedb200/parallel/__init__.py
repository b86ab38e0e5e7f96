"""Communication backends: in-process loopback (CPU plumbing), torch.distributed
NCCL/gloo (the baseline path) and the NVLink peer-memory runtime with fused collective kernels."""
from .placement import Placement, peers_per_rank, local_peer_ids, owner_of  # noqa: F401
from .comm import Comm, LoopbackComm, TorchDistComm, make_comm, init_distributed  # noqa: F401
