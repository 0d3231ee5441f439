"""vns-sop_b200: B200-native neighbourhood-evaluation engine for the VNS solver of the Set
Orienteering Problem (hot path of ctu-mrs/vns-sop): distance-table build, layered min-plus path
evaluation and batched move/exchange neighbourhoods as sm_100a CUDA kernels behind a C ABI
(include/vnsgpu.h).  The Python here is host glue (ctypes binding, instance files, sharding)."""
__version__ = "0.1.0"
