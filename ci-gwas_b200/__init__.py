"""cusk-b200: B200-native (sm_100a) CI-GWAS skeleton hot path.

Host-side Python mirror of the reference's operator interface (``Skeleton``, ``hetcor_skeleton``,
``cu_corr_pearson_npn`` ...) over the C ABI of ``libcusk_b200.so`` (``include/cusk_b200.h``).
There is no CPU fallback: every compute call raises when the CUDA library or a device is missing.
"""
from .lib import (  # noqa: F401
    CuskError,
    Context,
    Stats,
    block_chr,
    build_library,
    comm_unique_id,
    local_group,
    hanning_smoothing,
    hetcor_threshold,
    library_path,
    load_library,
    merge_block_outputs,
    mps_path,
    reformat_cuskss_merged,
    threshold_array,
)
