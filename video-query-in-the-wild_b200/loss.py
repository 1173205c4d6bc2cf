"""`misc_utils/metric_learning_loss.py`-compatible triplet criterion over the fused CUDA kernels
(SURVEY.md 8-b4).  ``TripletLoss(margin, sampling_method)(batch, labels)`` keeps the reference's
signature (:329-379); ``sampling_method="batchhard"`` (an extension -- the reference's Sampler has
random / semihard / distance / npair, :104-111) mines inside the kernel, the reference's own
samplers run on the host exactly as in the reference and hand their triplets to the same
loss/backward kernel.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib as L


def pdist(A, eps=1e-4):
    """Sampler.pdist (:276-289) on the GPU, canonical arithmetic."""
    A = A.detach().float().contiguous()
    out = torch.empty(A.shape[0], A.shape[0], dtype=torch.float32, device=A.device)
    L.check(L.lib().vq_pdist(L.ptr(A), C.c_int64(A.shape[0]), C.c_int(A.shape[1]), C.c_float(eps), L.ptr(out),
                             L.stream_of(A)))
    return out


class _BatchHardFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, batch, labels, margin):
        x = batch.detach().float().contiguous()
        grouped = x.dim() == 3
        G, B, d = (x.shape if grouped else (1,) + tuple(x.shape))
        lab = labels.to(device=x.device, dtype=torch.int32).contiguous()
        loss = torch.empty(G, dtype=torch.float32, device=x.device)
        grad = torch.empty_like(x)
        trip = torch.empty((G, B, 3), dtype=torch.int32, device=x.device)
        L.check(L.lib().vq_triplet_batchhard(L.ptr(x), L.ptr(lab), C.c_int64(G), C.c_int64(B), C.c_int(d),
                                             C.c_float(margin), L.ptr(loss), L.ptr(grad), L.ptr(trip),
                                             L.stream_of(x)))
        ctx.save_for_backward(grad)
        ctx.grouped = grouped
        ctx.mark_non_differentiable(trip)
        return (loss if grouped else loss[0]), (trip if grouped else trip[0])

    @staticmethod
    def backward(ctx, gloss, _gtrip):
        (grad,) = ctx.saved_tensors
        g = gloss.reshape(-1, 1, 1) * grad if ctx.grouped else gloss * grad
        return g, None, None


class _GivenFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, batch, trip, margin):
        x = batch.detach().float().contiguous()
        t = trip.to(device=x.device, dtype=torch.int32).contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        grad = torch.empty_like(x)
        L.check(L.lib().vq_triplet_given(L.ptr(x), C.c_int64(x.shape[0]), C.c_int(x.shape[1]), L.ptr(t),
                                         C.c_int64(t.shape[0]), C.c_float(margin), L.ptr(loss), L.ptr(grad),
                                         L.stream_of(x)))
        ctx.save_for_backward(grad)
        return loss[0]

    @staticmethod
    def backward(ctx, gloss):
        (grad,) = ctx.saved_tensors
        return gloss * grad, None, None


class Sampler:
    """Host samplers with the reference's selection rules (:113-196); distances come from the GPU
    pdist kernel.  They draw from numpy's global RNG exactly like the reference."""

    def __init__(self, method="random"):
        self.method = method
        self.give = {"semihard": self.semihardsampling, "random": self.randomsampling,
                     "batchhard": self.batchhardsampling}[method]

    pdist = staticmethod(pdist)

    @staticmethod
    def _np_labels(labels):
        return labels.detach().cpu().numpy() if isinstance(labels, torch.Tensor) else np.asarray(labels)

    def randomsampling(self, batch, labels):  # :113-148
        import itertools as it
        import random
        labels = self._np_labels(labels)
        unique_classes = np.unique(labels)
        indices = np.arange(len(batch))
        class_dict = {i: indices[labels == i] for i in unique_classes}
        trips = [list(it.product([x], [x], [y for y in unique_classes if x != y])) for x in unique_classes]
        trips = [x for y in trips for x in y]
        trips = [[x for x in list(it.product(*[class_dict[j] for j in i])) if x[0] != x[1]] for i in trips]
        trips = [x for y in trips for x in y]
        return random.sample(trips, batch.shape[0])

    def semihardsampling(self, batch, labels):  # :150-196
        labels = self._np_labels(labels)
        d_all = pdist(batch).cpu().numpy()
        out = []
        for i in range(batch.size(0)):
            l, d = labels[i], d_all[i]
            neg, pos = labels != l, labels == l
            pos[i] = False
            neg_mask = np.logical_and(neg, d < d[np.where(pos)[0]].max())
            pos_mask = np.logical_and(pos, d > d[np.where(neg)[0]].min())
            p = np.random.choice(np.where(pos_mask)[0]) if pos_mask.sum() > 0 else np.random.choice(np.where(pos)[0])
            n = np.random.choice(np.where(neg_mask)[0]) if neg_mask.sum() > 0 else np.random.choice(np.where(neg)[0])
            out.append([i, p, n])
        return out

    def batchhardsampling(self, batch, labels):
        lab = torch.as_tensor(self._np_labels(labels))
        _, trip = _BatchHardFn.apply(batch, lab, 0.0)
        return [t for t in trip.cpu().tolist() if t[0] >= 0]


class TripletLoss(torch.nn.Module):
    """:329-379.  forward(batch [BS,D] (or [G,BS,D] with batchhard), labels [BS]) -> mean triplet loss."""

    def __init__(self, margin=1, sampling_method="random"):
        super().__init__()
        self.margin = margin
        self.sampling_method = sampling_method
        self.sampler = Sampler(method=sampling_method)

    def forward(self, batch, labels):
        if self.sampling_method == "batchhard":
            lab = labels if isinstance(labels, torch.Tensor) else torch.as_tensor(np.asarray(labels))
            loss, _ = _BatchHardFn.apply(batch, lab, float(self.margin))
            return loss
        trip = torch.as_tensor(np.asarray(self.sampler.give(batch, labels), dtype=np.int32))
        return _GivenFn.apply(batch, trip, float(self.margin))


def loss_select(loss, opt, to_optim):
    """:33-83, triplet branch (the criteria north_star names).  Other criteria of the module are
    out of scope for the CUDA path."""
    if loss == "triplet":
        return TripletLoss(margin=opt.margin, sampling_method=opt.sampling), to_optim
    raise NotImplementedError("loss %r is outside the accelerated path (SURVEY.md 8-a12)" % (loss,))
