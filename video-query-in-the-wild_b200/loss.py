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
                     "distance": self.distanceweightedsampling, "npair": self.npairsampling,
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

    def inverse_sphere_distances(self, batch, dist, labels, anchor_label=None):
        """:291-324 for EVERY anchor in one launch: dist [B,B] clamped pdist -> q [B,B] (row i = the sampling
        distribution of anchor i).  ``anchor_label`` is accepted for signature compatibility and unused."""
        lab = torch.as_tensor(self._np_labels(labels)).to(device=dist.device, dtype=torch.int32).contiguous()
        dist = dist.detach().float().contiguous()
        q = torch.empty_like(dist)
        L.check(L.lib().vq_sphere_weights(L.ptr(dist), L.ptr(lab), C.c_int64(dist.shape[0]), C.c_int(batch.shape[-1]),
                                          L.ptr(q), L.stream_of(dist)))
        return q

    def distanceweightedsampling(self, batch, labels, lower_cutoff=0.5, upper_cutoff=1.4):  # :198-238
        labels = self._np_labels(labels)
        bs = batch.shape[0]
        distances = pdist(batch).clamp(min=lower_cutoff)
        q_all = self.inverse_sphere_distances(batch, distances, labels).cpu().numpy().astype(np.float64)
        positives, negatives = [], []
        for i in range(bs):
            pos = labels == labels[i]
            q_d_inv = q_all[i] / q_all[i].sum()        # np.random.choice wants sum(p) == 1 to ~1e-8
            pos[i] = 0
            positives.append(np.random.choice(np.where(pos)[0]))
            negatives.append(np.random.choice(bs, p=q_d_inv))
        return [[a, p, n] for a, p, n in zip(list(range(bs)), positives, negatives)]

    def npairsampling(self, batch, labels):  # :240-274
        labels = self._np_labels(labels)
        label_set, count = np.unique(labels, return_counts=True)
        label_set = label_set[count >= 2]
        pos_pairs = np.array([np.random.choice(np.where(labels == x)[0], 2, replace=False) for x in label_set])
        neg_tuples = []
        for idx in range(len(pos_pairs)):
            neg_tuples.append(pos_pairs[np.delete(np.arange(len(pos_pairs)), idx), 1])
        neg_tuples = np.array(neg_tuples)
        return [[a, p, *list(neg)] for (a, p), neg in zip(pos_pairs, neg_tuples)]

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


class _MarginFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, batch, beta_t, trip, margin):
        x = batch.detach().float().contiguous()
        t = trip.to(device=x.device, dtype=torch.int32).contiguous()
        be = beta_t.detach().to(device=x.device, dtype=torch.float32).contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        grad, gbeta = torch.empty_like(x), torch.empty_like(be)
        L.check(L.lib().vq_margin_given(L.ptr(x), C.c_int64(x.shape[0]), C.c_int(x.shape[1]), L.ptr(t),
                                        C.c_int64(t.shape[0]), L.ptr(be), C.c_float(margin), L.ptr(loss), L.ptr(grad),
                                        L.ptr(gbeta), C.c_void_p(0), L.stream_of(x)))
        ctx.save_for_backward(grad, gbeta)
        return loss[0]

    @staticmethod
    def backward(ctx, gloss):
        grad, gbeta = ctx.saved_tensors
        return gloss * grad, gloss * gbeta, None, None


class MarginLoss(torch.nn.Module):
    """:462-569.  Same constructor and forward(batch, labels); the per-triplet distances, hinges, pair count and
    both gradients (embeddings and the class margins beta) come from one C-ABI call."""

    def __init__(self, margin=0.2, nu=0, beta=1.2, n_classes=100, beta_constant=False, sampling_method="distance"):
        super().__init__()
        self.margin = margin
        self.n_classes = n_classes
        self.beta_constant = beta_constant
        self.beta_val = beta
        self.beta = beta if beta_constant else torch.nn.Parameter(torch.ones(n_classes) * beta)
        self.nu = nu
        self.sampling_method = sampling_method
        self.sampler = Sampler(method=sampling_method)

    def forward(self, batch, labels):
        if isinstance(labels, torch.Tensor):
            labels = labels.detach().cpu().numpy()
        sampled_triplets = self.sampler.give(batch, labels)
        trip = torch.as_tensor(np.asarray(sampled_triplets, dtype=np.int32))
        if self.beta_constant:
            beta_t = torch.full((trip.shape[0],), float(self.beta), device=batch.device)
        else:                                                                    # :543-547
            idx = torch.as_tensor(np.asarray(labels)[trip[:, 0].numpy()], dtype=torch.long, device=self.beta.device)
            beta_t = self.beta[idx].to(batch.device)
        loss = _MarginFn.apply(batch, beta_t, trip, float(self.margin))
        if self.nu:  # the reference adds an undefined name here (:563-564) and raises NameError
            raise NameError("name 'beta_regularisation_loss' is not defined")
        return loss


class _NPairFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, batch, tuples, l2):
        x = batch.detach().float().contiguous()
        t = tuples.to(device=x.device, dtype=torch.int32).contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        grad = torch.empty_like(x)
        L.check(L.lib().vq_npair_given(L.ptr(x), C.c_int64(x.shape[0]), C.c_int(x.shape[1]), L.ptr(t),
                                       C.c_int64(t.shape[0]), C.c_int(t.shape[1] - 2), C.c_float(l2), L.ptr(loss),
                                       L.ptr(grad), L.stream_of(x)))
        ctx.save_for_backward(grad)
        return loss[0]

    @staticmethod
    def backward(ctx, gloss):
        (grad,) = ctx.saved_tensors
        return gloss * grad, None, None


class NPairLoss(torch.nn.Module):
    """:384-457."""

    def __init__(self, l2=0.02):
        super().__init__()
        self.sampler = Sampler(method="npair")
        self.l2 = l2

    def forward(self, batch, labels):
        tuples = torch.as_tensor(np.asarray(self.sampler.give(batch, labels), dtype=np.int32))
        return _NPairFn.apply(batch, tuples, float(self.l2))


class _ProxyNCAFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, batch, proxies, labels):
        x = batch.detach().float().contiguous()
        pr = proxies.detach().to(device=x.device, dtype=torch.float32).contiguous()
        lab = labels.to(device=x.device, dtype=torch.int32).contiguous()
        loss = torch.empty(1, dtype=torch.float32, device=x.device)
        gx, gp = torch.empty_like(x), torch.empty_like(pr)
        L.check(L.lib().vq_proxynca(L.ptr(x), L.ptr(pr), L.ptr(lab), C.c_int64(x.shape[0]), C.c_int64(pr.shape[0]),
                                    C.c_int(x.shape[1]), L.ptr(loss), L.ptr(gx), L.ptr(gp), L.stream_of(x)))
        ctx.save_for_backward(gx, gp)
        return loss[0]

    @staticmethod
    def backward(ctx, gloss):
        gx, gp = ctx.saved_tensors
        return gloss * gx, gloss * gp, None


class ProxyNCALoss(torch.nn.Module):
    """:574-635."""

    def __init__(self, num_proxies, embedding_dim):
        super().__init__()
        self.num_proxies = num_proxies
        self.embedding_dim = embedding_dim
        self.PROXIES = torch.nn.Parameter(torch.randn(num_proxies, self.embedding_dim) / 8)
        self.all_classes = torch.arange(num_proxies)

    def forward(self, batch, labels):
        lab = labels if isinstance(labels, torch.Tensor) else torch.as_tensor(np.asarray(labels))
        return _ProxyNCAFn.apply(batch, self.PROXIES, lab)


class CEClassLoss(torch.nn.Module):
    """:641-667 -- a Linear layer + CrossEntropyLoss; no pairwise distances, so it stays plain torch (library ops)."""

    def __init__(self, inp_dim, n_classes):
        super().__init__()
        self.mapper = torch.nn.Sequential(torch.nn.Linear(inp_dim, n_classes))
        self.ce_loss = torch.nn.CrossEntropyLoss()

    def forward(self, batch, labels):
        lab = labels if isinstance(labels, torch.Tensor) else torch.as_tensor(np.asarray(labels))
        return self.ce_loss(self.mapper(batch), lab.to(device=batch.device, dtype=torch.long))


def loss_select(loss, opt, to_optim):
    """:33-83 -- same names, parameters and optimiser groups as the reference."""
    if loss == "triplet":
        criterion = TripletLoss(margin=opt.margin, sampling_method=opt.sampling)
    elif loss == "npair":
        criterion = NPairLoss(l2=opt.l2npair)
    elif loss == "marginloss":
        criterion = MarginLoss(margin=opt.margin, nu=opt.nu, beta=opt.beta, n_classes=opt.nclass, sampling_method=opt.sampling)
        to_optim += [{"params": criterion.parameters(), "lr": opt.beta_lr, "weight_decay": 0}]
    elif loss == "proxynca":
        criterion = ProxyNCALoss(num_proxies=opt.nclass,
                                 embedding_dim=opt.classembed if "num_cluster" in vars(opt).keys() else opt.embed_dim)
        to_optim += [{"params": criterion.parameters(), "lr": opt.proxy_lr}]
    elif loss == "crossentropy":
        criterion = CEClassLoss(n_classes=opt.nclass, inp_dim=opt.embed_dim)
        to_optim += [{"params": criterion.parameters(), "lr": opt.lr, "weight_decay": 0}]
    else:
        raise Exception("Loss {} not available!".format(loss))
    return criterion, to_optim


class _MemoryLogitsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mem, temperature):
        xx = x.detach().float().contiguous()
        mm = mem.detach().to(device=xx.device, dtype=torch.float32).contiguous()
        logits = torch.empty((xx.shape[0], mm.shape[0]), dtype=torch.float32, device=xx.device)
        L.check(L.lib().vq_memory_logits(L.ptr(xx), L.ptr(mm), C.c_int64(xx.shape[0]), C.c_int64(mm.shape[0]),
                                         C.c_int(xx.shape[1]), C.c_float(temperature), L.ptr(logits), C.c_void_p(0),
                                         C.c_void_p(0), C.c_void_p(0), L.stream_of(xx)))
        ctx.save_for_backward(xx, mm)
        ctx.temperature = temperature
        return logits

    @staticmethod
    def backward(ctx, g):
        xx, mm = ctx.saved_tensors
        g = g.float().contiguous()
        gx = torch.empty_like(xx) if ctx.needs_input_grad[0] else None
        gm = torch.empty_like(mm) if ctx.needs_input_grad[1] else None
        L.check(L.lib().vq_memory_logits(L.ptr(xx), L.ptr(mm), C.c_int64(xx.shape[0]), C.c_int64(mm.shape[0]),
                                         C.c_int(xx.shape[1]), C.c_float(ctx.temperature), C.c_void_p(0), L.ptr(g),
                                         L.ptr(gx), L.ptr(gm), L.stream_of(xx)))
        return gx, gm, None


def memory_logits(x, memory, temperature):
    """The VA / VASA heads' memory-distance logits (models/resnet18_va.py:172-184, resnet18_vasa.py:198-206,
    :217-229): ``-torch.norm(F.normalize(x_b) - memory, p=2, dim=1) / temperature`` for every b, one call instead
    of the reference's Python loop over the batch; differentiable w.r.t. ``x`` (and ``memory`` if it requires grad)."""
    return _MemoryLogitsFn.apply(x, memory, float(temperature))
