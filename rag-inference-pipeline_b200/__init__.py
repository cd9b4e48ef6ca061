"""B200-native exact flat inner-product retrieval (drop-in for the reference's FAISS hot path).

Public surface:
  B200FlatIPStore            component with the reference FAISSStore's interface (store.py)
  create_b200_flat_ip_store  factory with the reference's (settings, config) signature
  install                    registers the factory under "faiss" / "faiss_store"
  FlatIPIndex                faiss.Index-like wrapper over the C ABI (one GPU, one shard)
  FlatIPGroup                the same surface over several GPUs of this process (fip_group_*: row shards,
                             one host thread + stream per device, K4 exchange + merge)
  IVFFlatL2Index             faiss.IndexIVFFlat-like wrapper (METRIC_L2, nprobe) over the C ABI
  CandidateFeed              batched id-matrix -> documents fetch for the rerank stage (DocumentStore interface)
  ShardedFlatIPSearcher      row-sharded search across torch.distributed ranks (K4 over peer memory, or NCCL + K3 merge)
"""

from .candidate_feed import CandidateDocument, CandidateFeed, create_candidate_feed
from .factory import COMPONENT_TYPE, create_b200_flat_ip_store, install
from .group import FlatIPGroup, calibrate_devices
from .index import FlatIPIndex, merge_device, merge_strided_device
from .ivf import IVFFlatL2Index
from .store import B200FlatIPStore

__all__ = [
    "B200FlatIPStore", "CandidateDocument", "CandidateFeed", "create_candidate_feed", "FlatIPIndex", "FlatIPGroup", "calibrate_devices", "IVFFlatL2Index", "ShardedFlatIPSearcher", "COMPONENT_TYPE",
    "create_b200_flat_ip_store", "install", "merge_device", "merge_strided_device",
]


def __getattr__(name):
    if name == "ShardedFlatIPSearcher":  # torch is only imported when the sharded path is used
        from .sharded import ShardedFlatIPSearcher
        return ShardedFlatIPSearcher
    raise AttributeError(name)
