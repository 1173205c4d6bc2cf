"""CPU oracle (numpy, float64 internally) for the remaining metric-learning criteria and the
memory-distance logits (8-f3).  TEST INFRASTRUCTURE ONLY: imported by tests/ (and nothing in the product).
Every function restates the reference lines it cites; pinned against reference_golden_losses.npz, which was
minted by running misc_utils/metric_learning_loss.py itself (tests/golden/make_golden_losses.py)."""
import numpy as np


def margin_given(emb, trip, beta_t, margin):
    """MarginLoss.forward on a given triplet list (metric_learning_loss.py:520-565), nu = 0.
    beta_t[t] = beta[labels[anchor_t]] (:543-547) or the constant.  Returns loss, grad_emb, grad_beta_t."""
    x = np.asarray(emb, np.float64)
    trip = np.asarray(trip, np.int64)
    beta_t = np.broadcast_to(np.asarray(beta_t, np.float64), (len(trip),))
    a, p, n = x[trip[:, 0]], x[trip[:, 1]], x[trip[:, 2]]
    d_ap = np.sqrt(((a - p) ** 2).sum(1) + 1e-8)                      # :527-533
    d_an = np.sqrt(((a - n) ** 2).sum(1) + 1e-8)                      # :534-540
    pos = np.maximum(d_ap - beta_t + margin, 0.0)                     # :549
    neg = np.maximum(beta_t - d_an + margin, 0.0)                     # :550
    # :552-554 `(pos_loss > 0.0) + (neg_loss > 0.0)` adds two BOOL tensors = logical or: triplets with an active
    # term are counted once (not the number of active terms)
    pc = float(((pos > 0) | (neg > 0)).sum())
    den = pc if pc != 0.0 else 1.0
    loss = (pos + neg).sum() / den                                    # :557-561
    g = np.zeros_like(x)
    kap = (pos > 0) / (den * d_ap)
    kan = (neg > 0) / (den * d_an)
    np.add.at(g, trip[:, 0], kap[:, None] * (a - p) - kan[:, None] * (a - n))
    np.add.at(g, trip[:, 1], -kap[:, None] * (a - p))
    np.add.at(g, trip[:, 2], kan[:, None] * (a - n))
    gbeta = ((neg > 0).astype(np.float64) - (pos > 0)) / den
    return loss, g, gbeta


def npair_given(emb, tuples, l2):
    """NPairLoss.forward on given tuples [P, 2+N] = (anchor, positive, negatives...) (:397-457)."""
    x = np.asarray(emb, np.float64)
    tuples = np.asarray(tuples, np.int64)
    P = len(tuples)
    g = np.zeros_like(x)
    total = 0.0
    for t in tuples:
        a, p, neg = x[t[0]], x[t[1]], x[t[2:]]
        s = (neg - p) @ a                                             # :409 anchor.mm((negatives-positive).T)
        ex = np.exp(s)
        total += np.log(1.0 + ex.sum())                               # :407-411
        pi = ex / (1.0 + ex.sum())
        g[t[0]] += (pi[:, None] * (neg - p)).sum(0) / P
        g[t[1]] += -pi.sum() * a / P
        np.add.at(g, t[2:], pi[:, None] * a[None, :] / P)
    ws = np.array([(x[t[0]] ** 2 + x[t[1]] ** 2).sum() for t in tuples])  # :413-423
    loss = total / P + l2 * ws.mean()                                 # :446-457 (mean of loss_i + l2*mean(ws))
    for t in tuples:
        g[t[0]] += l2 * 2.0 * x[t[0]] / P
        g[t[1]] += l2 * 2.0 * x[t[1]] / P
    return loss, g


def _normalize3(x):
    nrm = np.maximum(np.sqrt((x ** 2).sum(1, keepdims=True)), 1e-12)  # F.normalize eps
    return 3.0 * x / nrm, nrm


def _normalize3_bwd(x, nrm, g):
    xh = x / nrm
    return 3.0 / nrm * (g - xh * (xh * g).sum(1, keepdims=True))


def proxynca(emb, proxies, labels):
    """ProxyNCALoss.forward (:593-635): batch and proxies scaled to norm 3; loss = mean_b( d(x_b, p_y) +
    logsumexp_{c != y}(-d(x_b, p_c)) ) with d = squared L2.  Returns loss, grad_emb, grad_proxies."""
    x = np.asarray(emb, np.float64)
    pr = np.asarray(proxies, np.float64)
    y = np.asarray(labels, np.int64)
    B, C = len(x), len(pr)
    bn, bnrm = _normalize3(x)                                         # :602
    pn, pnrm = _normalize3(pr)                                        # :603
    d2 = ((bn[:, None, :] - pn[None, :, :]) ** 2).sum(-1)             # :620-625
    mask = np.ones((B, C), bool)
    mask[np.arange(B), y] = False
    z = np.where(mask, -d2, -np.inf)
    m = z.max(1, keepdims=True)
    lse = m[:, 0] + np.log(np.exp(z - m).sum(1))
    loss = (d2[np.arange(B), y] + lse).mean()                         # :627-630
    sm = np.exp(z - lse[:, None])                                     # softmax over the negatives
    K = -sm
    K[np.arange(B), y] = 1.0
    W = 2.0 * K / B                                                   # dL/d d2 * 2
    diff = bn[:, None, :] - pn[None, :, :]
    gbn = (W[:, :, None] * diff).sum(1)
    gpn = -(W[:, :, None] * diff).sum(0)
    return loss, _normalize3_bwd(x, bnrm, gbn), _normalize3_bwd(pr, pnrm, gpn)


def memory_logits(x, mem, temperature, w=None):
    """reg_logits[b] = -||x_hat_b - M||_2 / temperature with x_hat = F.normalize(x) (models/resnet18_va.py:172-184,
    resnet18_vasa.py:198-206; word_logits :217-229 is the same op on the adaptor output and the semantic memory).
    With w: also d(sum(logits*w))/dx (norm's subgradient at distance 0 is 0, as torch)."""
    x = np.asarray(x, np.float64)
    M = np.asarray(mem, np.float64)
    nrm = np.maximum(np.sqrt((x ** 2).sum(1, keepdims=True)), 1e-12)
    xh = x / nrm
    diff = xh[:, None, :] - M[None, :, :]
    dist = np.sqrt((diff ** 2).sum(-1))
    logits = -dist / temperature
    if w is None:
        return logits
    coef = np.where(dist > 0, -np.asarray(w, np.float64) / temperature / np.where(dist > 0, dist, 1.0), 0.0)
    gxh = (coef[:, :, None] * diff).sum(1)
    gx = (gxh - xh * (xh * gxh).sum(1, keepdims=True)) / nrm
    return logits, gx


def inverse_sphere_distances(dist_row, labels, anchor_label, dim):
    """Sampler.inverse_sphere_distances (:291-324) for one anchor; dist_row already clamped at the lower cutoff."""
    d = np.asarray(dist_row, np.float64)
    lab = np.asarray(labels)
    log_q = (2.0 - float(dim)) * np.log(d) - (float(dim - 3) / 2) * np.log(1.0 - 0.25 * d ** 2)
    same = lab == anchor_label
    log_q[same] = 0
    q = np.exp(log_q - log_q.max())
    q[same] = 0
    return q / q.sum()
