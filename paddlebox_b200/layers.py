"""One namespace with the layer names a PaddleBox script uses.

``fluid.layers`` (python/paddle/fluid/layers/nn.py:714-895, :3490, :14947) and
``fluid.contrib.layers`` (python/paddle/fluid/contrib/layers/nn.py:1497-2896) spread the CTR layers
over two modules; a ported script can ``from paddlebox_b200 import layers`` and keep its calls.  The
tensors differ from the static-graph originals in one way only: variable-length slot inputs arrive as
one slot-major tensor plus a ``lod`` (``ops.pull_box_sparse.SlotBatch``) instead of a list of LoDTensors.
"""
from __future__ import annotations

from .ops.ctr_ops import (BatchFC, CrossNormHadamard, batch_fc, build_rank_offset,  # noqa: F401
                          cross_norm_layer_hadamard, filter_by_instag, fused_concat, fused_seq_tensor,
                          fused_seqpool_concat, fused_seqpool_cvm_with_pcoc, partial_concat, partial_sum,
                          rank_attention, rank_attention2, scaled_fc, scaled_int8fc, shuffle_batch)
from .ops.functional import (DataNorm, continuous_value_model, ctr_head, data_norm, fm_interaction,  # noqa: F401
                             masked_data_norm, sigmoid_log_loss, store_q_value, stored_q_values)
from .ops.misc_layers import (fused_elemwise_activation, fused_embedding_seq_pool,  # noqa: F401
                              pow2_decay_with_linear_warmup, sparse_embedding)
from .ops.pull_box_sparse import (SlotBatch, pull_box_extended_sparse, pull_box_sparse, sparse_embedding_pool,  # noqa: F401
                                  split_by_slot)
from .ops.seqpool import (SeqpoolAttrs, fused_seqpool_cvm, fused_seqpool_cvm_tradew, fused_seqpool_cvm_with_conv,  # noqa: F401
                          fused_seqpool_cvm_with_credit, fused_seqpool_cvm_with_diff_thres, split_slots)
from .ops.sequence import (bpr_loss, hash, log_loss, margin_rank_loss, match_matrix_tensor, rank_loss,  # noqa: F401
                           sequence_concat, sequence_conv, sequence_enumerate, sequence_expand, sequence_expand_as,
                           sequence_first_step, sequence_last_step, sequence_mask, sequence_pad, sequence_pool,
                           sequence_reshape, sequence_reverse, sequence_scatter, sequence_slice, sequence_softmax,
                           sequence_topk_avg_pooling, sequence_unpad, teacher_student_sigmoid_loss)
from .metrics import Auc, auc  # noqa: F401
from .ops.tdm import TDMTree, tdm_child, tdm_sampler  # noqa: F401
from .ps_gpu import pull_gpups_sparse  # noqa: F401

# the reference's underscore-prefixed spellings
_pull_box_sparse = pull_box_sparse
_pull_box_extended_sparse = pull_box_extended_sparse
_pull_gpups_sparse = pull_gpups_sparse
_store_q_value = store_q_value


def pull_cache_value(cache, idx):
    """``pull_cache_value`` op: rows of the per-pass replica cache (``side_tables.GpuReplicaCache``)."""
    return cache.pull_cache_value(idx)


def lookup_input(table, idx):
    """``lookup_input`` op: rows of the ``side_tables.InputTable`` for offsets produced at parse time."""
    return table.lookup_input(idx)
