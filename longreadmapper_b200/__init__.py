"""longreadmapper_b200 -- B200 (sm_100a) implementation of AccSeqV9's per-read mapping hot path.

The product is the C-ABI library ``libaccseq_b200.so`` (include/accseq_b200.h) plus the plain-C host helpers
``libaccseq_host.so`` (include/accseq_host.h).  This package is only the ctypes view of both that the tests,
``bench.py`` and ``__graft_entry__`` use; there is no CPU fallback behind it.
"""
from . import capi, shard
from .capi import (COMPACT_DTYPE, pack_reads, results_from_compact, AS_F_GACT_BITPAR, AS_F_GACT_WARP, AS_F_MYERS, AS_F_KEEP_READS, AS_F_NO_ALIGN, AS_F_NO_LC2, AS_F_RLE_CIGAR, AS_F_WIDE_INDEX, RESULT_DTYPE, AccSeqError, GpuMapper, Params, build_native,
                   cat_image, cigar_text, cigar_text_rle, cuda_lib, device_count, host_lib, open_opts, sim_genome, sim_plant_repeats, sim_reads)

__all__ = ["COMPACT_DTYPE", "pack_reads", "results_from_compact", "AS_F_GACT_BITPAR", "AS_F_GACT_WARP", "AS_F_MYERS", "AS_F_KEEP_READS", "AS_F_NO_ALIGN", "AS_F_NO_LC2", "AS_F_RLE_CIGAR", "AS_F_WIDE_INDEX", "RESULT_DTYPE", "AccSeqError", "GpuMapper", "Params", "build_native", "cat_image", "cigar_text", "cigar_text_rle",
           "cuda_lib", "device_count", "host_lib", "open_opts", "sim_genome", "sim_plant_repeats", "sim_reads"]
