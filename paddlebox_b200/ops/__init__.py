"""Operators: sparse table, pull/push, fused_seqpool_cvm family, tcgen05 towers, CTR utilities."""
