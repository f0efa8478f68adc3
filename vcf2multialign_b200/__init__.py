"""vcf2multialign_b200 — B200 (sm_100a) implementation of vcf2multialign's aligned-sequence materialisation path.

Everything here runs through libv2m_cuda.so (include/v2m_cuda.h); there is no CPU fallback."""
from ._lib import PLOIDY_MAX, V2MError, lib  # noqa: F401
from .graph import EDGE_MAX, GT_NULL, DeviceGraph, Plan, VariantGraph, paths_from_genotypes, paths_from_genotypes_both  # noqa: F401
from .output import founder_sequence_greedy_output, haplotype_output, output_delegate, transpose_matrix  # noqa: F401
