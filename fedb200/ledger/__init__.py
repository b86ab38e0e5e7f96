"""Hash-chained ledger of weight updates (C24; kernels K6 ``sha256_send`` / K7 ``block_validate``).

The reference only *names* a blockchain ("BC-FL algorithms", README.md:10, and the plot title at
All_graphs_IMDB_dataset.ipynb:1048); no block structure, hash or validation exists there.  This
package defines one: every weight transfer appends a block whose ``update_root`` is the SHA-256
Merkle root of the bytes that were sent."""
from .merkle import merkle_root, leaf_digests, reduce_level, sha256  # noqa: F401
from .block import BlockHeader, HEADER_BYTES  # noqa: F401
from .chain import Ledger, LedgerError  # noqa: F401
