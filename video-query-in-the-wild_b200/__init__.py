"""vqwild_b200 -- B200-native (sm_100a) query->gallery matching path of
dongzhuoyao/video-query-in-the-wild: IndexFlatL2-compatible search, the evaluator's ranking /
mAP / R@N, temporal NMS and the in-batch triplet loss, as thin callers into libvqwild.so
(include/vqwild.h).  There is no CPU fallback."""
from ._lib import (PREC_AUTO, PREC_BF16, PREC_EXACT, PREC_SPLIT3, SO_PATH, VQ_BF16, VQ_F32, VqError, build,
                   lib)
from .index import IndexFlatL2, faiss_module
from .metrics import (Average, average_precision_batch, average_precision_score, evaluation_metric,
                      generate_multi_query, nms_cpu, nms_segmented, nms_segmented_static, recall_at_k, ap_at_k)
from .retrieval import ARV_Retrieval, ARV_Retrieval_Clip, ARV_Retrieval_Moment, wordembed_topk


def __getattr__(name):  # torch-dependent pieces are imported lazily
    if name in ("TripletLoss", "Sampler", "loss_select", "pdist", "MarginLoss", "NPairLoss", "ProxyNCALoss", "CEClassLoss",
                "memory_logits"):
        from . import loss
        return getattr(loss, name)
    if name in ("pool_normalize", "topk_merge", "topk_pack", "topk_merge_keys", "topk_merge_heads", "exchange_merge", "l2sq", "sharded_search",
                "window_means"):
        from . import ops
        return getattr(ops, name)
    raise AttributeError(name)
