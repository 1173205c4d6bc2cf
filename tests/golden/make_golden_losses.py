"""Golden vectors for the remaining criteria of misc_utils/metric_learning_loss.py and the
memory-distance logits of the VA / VASA heads (8-f3), minted by running the REFERENCE modules on CPU
(authoring container only):

    python tests/golden/make_golden_losses.py      # needs /root/reference; writes reference_golden_losses.npz

  * MarginLoss (:462-569) with sampling_method="distance" and "semihard", per-class and constant beta
  * NPairLoss (:384-457), ProxyNCALoss (:574-635)
  * Sampler.inverse_sphere_distances (:291-324), the sampled tuples of distanceweightedsampling / npairsampling
One shim: MarginLoss casts with `.type(torch.cuda.FloatTensor)` (:547, :554), which needs a CUDA device; the name
is pointed at torch.FloatTensor for this run (same cast, CPU storage).
The memory logits (models/resnet18_va.py:176-184, resnet18_vasa.py:198-206, :217-229) sit inside a network's
forward; their eight lines are re-typed below verbatim on seeded inputs ("restated", not reference output).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path.insert(0, REF)


def main():
    tc = types.ModuleType("termcolor")
    tc.colored = lambda s, *a, **k: s
    sys.modules["termcolor"] = tc
    from misc_utils import metric_learning_loss as mll
    torch.cuda.FloatTensor = torch.FloatTensor  # the shim described above
    out = {}
    rng = np.random.RandomState(5)
    B, D, NC = 48, 32, 12
    labels = np.repeat(np.arange(NC), B // NC)
    centre = rng.randn(NC, D).astype(np.float32)
    emb = rng.randn(B, D).astype(np.float32) * 0.7 + centre[labels]
    emb /= np.linalg.norm(emb, axis=1, keepdims=True)
    out["emb"], out["labels"] = emb, labels

    # ---- MarginLoss ----
    for tag, sampling, beta_constant in (("mg_dist", "distance", False), ("mg_semi", "semihard", True)):
        np.random.seed(7)
        crit = mll.MarginLoss(margin=0.2, nu=0, beta=1.2 if not beta_constant else 0.9, n_classes=NC,
                              beta_constant=beta_constant, sampling_method=sampling)
        if not beta_constant:
            with torch.no_grad():
                crit.beta += torch.linspace(-0.3, 0.3, NC)
            out[tag + "_beta"] = crit.beta.detach().numpy().copy()
        x = torch.from_numpy(emb.copy()).requires_grad_(True)
        state = np.random.get_state()
        trip = crit.sampler.give(x, labels)
        np.random.set_state(state)
        loss = crit(x, labels)
        loss.backward()
        out[tag + "_trip"] = np.array(trip, np.int32)
        out[tag + "_loss"], out[tag + "_grad"] = loss.item(), x.grad.numpy().copy()
        if not beta_constant:
            out[tag + "_gbeta"] = crit.beta.grad.numpy().copy()
    # the sampling weights themselves (anchor 0 and 17)
    samp = mll.Sampler("distance")
    xt = torch.from_numpy(emb)
    dist = samp.pdist(xt).clamp(min=0.5)
    out["sphere_q"] = np.stack([samp.inverse_sphere_distances(xt, dist[i], labels, labels[i]) for i in (0, 17)])

    # ---- NPairLoss ----
    np.random.seed(9)
    crit = mll.NPairLoss(l2=0.02)
    x = torch.from_numpy((emb * 1.7).copy()).requires_grad_(True)
    state = np.random.get_state()
    tuples = crit.sampler.give(x, labels)
    np.random.set_state(state)
    loss = crit(x, labels)
    loss.backward()
    out["np_scale"] = 1.7
    out["np_tuples"] = np.array(tuples, np.int32)
    out["np_loss"], out["np_grad"] = loss.item(), x.grad.numpy().copy()

    # ---- ProxyNCALoss ----
    torch.manual_seed(3)
    crit = mll.ProxyNCALoss(num_proxies=NC, embedding_dim=D)
    x = torch.from_numpy((emb * 2.5).copy()).requires_grad_(True)
    loss = crit(x, labels)
    loss.backward()
    out["px_scale"] = 2.5
    out["px_proxies"] = crit.PROXIES.detach().numpy().copy()
    out["px_loss"], out["px_grad"] = loss.item(), x.grad.numpy().copy()
    out["px_gproxies"] = crit.PROXIES.grad.numpy().copy()

    # ---- memory-distance logits (restated lines, see the header) ----
    import torch.nn.functional as F
    K, temperature = 20, 0.05
    mem = F.normalize(torch.from_numpy(rng.randn(K, D).astype(np.float32)), p=2, dim=1)
    mem[3] = F.normalize(torch.from_numpy(emb[5] * 3.0 + 0.05 * emb[6]), p=2, dim=0)   # a memory slot close to an embedding
    cls_embed = torch.from_numpy((emb[:16] * 3.0).copy()).requires_grad_(True)
    normalized_cls_embed = F.normalize(cls_embed, p=2, dim=-1)
    reg_logits = torch.ones([16, K])
    for b in range(16):                                                  # resnet18_va.py:178-184
        tmp = -torch.norm(normalized_cls_embed[b] - mem, p=2, dim=1) / temperature
        reg_logits[b] = tmp
    w = torch.from_numpy(rng.randn(16, K).astype(np.float32))
    (reg_logits * w).sum().backward()
    out["ml_x"], out["ml_mem"], out["ml_w"], out["ml_temp"] = cls_embed.detach().numpy().copy(), mem.numpy().copy(), w.numpy(), temperature
    out["ml_logits"], out["ml_gx"] = reg_logits.detach().numpy().copy(), cls_embed.grad.numpy().copy()

    np.savez_compressed(os.path.join(HERE, "reference_golden_losses.npz"), **out)
    for k, v in out.items():
        print(k, getattr(v, "shape", v))


if __name__ == "__main__":
    main()
