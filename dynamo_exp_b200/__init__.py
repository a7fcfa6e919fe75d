"""dynamo_exp_b200 — B200-native KV-cache data plane (drop-in for the reference's kv::layer / kv::storage / lib/tokens path).

The package is a thin host layer over the C-ABI shared library `libkvdp.so` (CUDA, sm_100a). Importing the package
does not load the library; the first operation does, and raises if it is missing (there is no CPU fallback).
"""
from . import _lib  # noqa: F401
from .storage import (CudaPinnedMemory, DeviceStorageFromAny, DeviceStorageOwned, DType, KvError, OwnedStorage,  # noqa: F401
                      Storage, StorageType, SystemMemory, TensorView)
from .layer import (CopyStream, CopyStreamBlockMap, KvBlockDetails, KvBlockStorage, KvLayer, KvLayout,  # noqa: F401
                    KvModelDetails)
from .tokens import (PartialTokenBlock, TokenBlock, TokenBlockSequence, Tokens, TwinTokens, XXH3_SEED,  # noqa: F401
                     compute_block_hash_for_seq, compute_hash, hash_token_blocks)
from .publisher import KvEventPublisher  # noqa: F401
from .transfer import CopyPlan, HostTier, KvContiguousBlocks, TransferEngine  # noqa: F401

__version__ = "0.2.0"
