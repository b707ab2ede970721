"""charon_b200 — B200-native read-classification hot path of Charon (`charon dehost` / `charon classify`).

csrc/   hand-written sm_100a kernels + the C-ABI (include/charon_gpu.h) -> libcharon_gpu.so
host/   C++14 host program: index loader, FASTA/Q ingest, pre-filters, TSV writer, CLI
api.py  ctypes mirror used by the Python tests and bench.py
"""
from .api import (GpuClassifier, DehostOptions, ClassifyOptions, CharonGpuError, load_library, LIB_PATH, PackedBatch, pack_batch,  # noqa: F401
                  BASES_ASCII, BASES_RANK4, BASES_2BIT)
