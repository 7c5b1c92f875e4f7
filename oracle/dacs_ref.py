"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the DACS heads/losses and FedAvg.

Restates federated/src/models.py:61-260 (GRL, gumbel_softmax, CosFace, RecallLoss) and the
forward at :375-631, plus federated/src/utils.py:240-249 (average_weights).  The Gumbel noise
is an explicit argument (the reference draws `-log(Exp(1))` from torch's RNG stream,
models.py:111-116) so that the oracle, the reference (monkey-patched in make_golden.py) and
the CUDA kernel see identical noise.
"""
from __future__ import annotations

import copy
from types import SimpleNamespace
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import data2vec_ref as R


class _GRL(torch.autograd.Function):
    """federated/src/models.py:61-73."""

    @staticmethod
    def forward(ctx, x, lam):
        ctx.lam = float(lam)
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return -ctx.lam * g, None


def gumbel_hard_mask(score2: torch.Tensor, gumbels: torch.Tensor, tau: float, hard0: Optional[torch.Tensor] = None) -> torch.Tensor:
    """federated/src/models.py:75-129 with hard=True, then `[..., 0]`.
    score2, gumbels: [..., 2].  hard0 (bool [...]): take these hard decisions instead of the argmax (parity "given the
    masks": the decisions sit behind a comparison that another implementation's rounding may flip at near-ties); the
    straight-through gradient still flows through y_soft."""
    y_soft = ((score2 + gumbels) / tau).softmax(-1)
    if hard0 is not None:
        h0 = hard0.to(y_soft.dtype)
        return h0 - y_soft[..., 0].detach() + y_soft[..., 0]
    index = y_soft.max(-1, keepdim=True)[1]
    y_hard = torch.zeros_like(score2).scatter_(-1, index, 1.0)
    return (y_hard - y_soft.detach() + y_soft)[..., 0]


def recall_loss(logits2: torch.Tensor, target: torch.Tensor, weight, kind: str = "recall") -> torch.Tensor:
    """federated/src/models.py:187-260 on [N, C] logits (last_size == 1)."""
    w = torch.tensor(weight, dtype=torch.float32, device=logits2.device)
    w = w / w.sum()
    x = logits2.float()
    N, C = x.shape
    pt = F.log_softmax(x, dim=1).exp()
    onehot = torch.zeros(N, C, device=x.device).scatter_(1, target.view(N, 1).long(), 1.0)
    tp = pt * onehot
    recall = (tp + 1e-5) / (onehot + 1e-5)
    precision = (tp + 1e-5) / (pt + 1e-5)
    f1 = 2 * recall * precision / (recall + precision)
    recall_ori = recall * w * C
    precision_ori = precision * w * C
    f1 = f1 * w * C
    recall_c = (1.0 - recall) * w * C
    precision_c = (1.0 - precision) * w * C
    return {"recall": recall_c.mean(), "prec": precision_c.mean(), "f1": 1 - f1.mean(),
            "recall_ori": 1 - recall_ori.mean(), "prec_ori": 1 - precision_ori.mean()}[kind]


def cosface_loss(x: torch.Tensor, labels: torch.Tensor, fc_weight: torch.Tensor, s: float = 30.0, m: float = 0.4):
    """federated/src/models.py:131-185, loss_type='cosface'.  The weight-normalise loop there
    (:167-168) rebinds a local and has no effect, so `fc_weight` is used as stored.  Closed form of
    the N x N diagonal gather (:172) and the per-row Python loop (:178)."""
    xn = F.normalize(x, p=2, dim=1)
    wf = xn @ fc_weight.t()
    idx = labels.view(-1, 1).long()
    tgt = wf.gather(1, idx).squeeze(1)
    oth = wf.gather(1, 1 - idx).squeeze(1)
    num = s * (tgt - m)
    L = num - torch.log(torch.exp(num) + torch.exp(s * oth))
    return -L.mean()


def dacs_forward(sd, cfg: R.ArchConfig, args, input_values, attention_mask, labels, dementia_labels,
                 gumbels_lm: torch.Tensor, gumbels_ad: torch.Tensor, hidden: Optional[torch.Tensor] = None,
                 hard_lm: Optional[torch.Tensor] = None, hard_ad: Optional[torch.Tensor] = None):
    """federated/src/models.py:375-631 with dropout = 0.  Returns a dict with every loss term, the
    stage-selected `loss` and the returned (lm-masked) `logits`.  `hidden` may be supplied to skip
    the encoder (used when only the heads are under test)."""
    d = cfg.hidden_size
    if hidden is None:
        hidden, _ = R.base_model(sd, cfg, input_values, attention_mask)
    h = hidden
    lam = float(args.LAMBDA)
    ad_unmask = F.linear(h, sd["dementia_head.weight"], sd["dementia_head.bias"])
    score = F.linear(h, sd["arbitrator.weight"], sd["arbitrator.bias"])
    lm_score = torch.stack((score[..., :d], score[..., d:2 * d]), -1)
    ad_score = torch.stack((score[..., 2 * d:3 * d], score[..., 3 * d:]), -1)
    if args.TOGGLE_RATIO != 0:
        r = args.TOGGLE_RATIO
        lm_score = torch.stack(((lm_score[..., 1] - lm_score[..., 0]) * r + lm_score[..., 0], lm_score[..., 1]), -1)
        ad_score = torch.stack(((ad_score[..., 1] - ad_score[..., 0]) * r + ad_score[..., 0], ad_score[..., 1]), -1)
    lm_mask = gumbel_hard_mask(lm_score, gumbels_lm, args.GS_TAU, hard_lm)
    ad_mask = gumbel_hard_mask(ad_score, gumbels_ad, args.GS_TAU, hard_ad)
    lm_masked = lm_mask * h
    ad_masked = ad_mask * h
    lin = lambda x, n: F.linear(x, sd[n + ".weight"], sd[n + ".bias"])
    logits_unmask = lin(h, "lm_head")
    logits = lin(lm_masked, "lm_head")
    ad_lm = lin(lm_masked, "dementia_head")
    ad_mean_r = _GRL.apply(ad_lm.mean(1), lam)
    ad_mean_unmask = ad_unmask.mean(1)
    logits_r = lin(ad_masked, "lm_head")
    ad_mean = lin(ad_masked, "dementia_head").mean(1)
    out = dict(logits=logits, logits_unmask=logits_unmask, logits_r=logits_r, lm_mask=lm_mask, ad_mask=ad_mask,
               hidden=h, loss=None)
    if labels is None:
        return out
    il, tl, ft = R.ctc_inputs(cfg, input_values, attention_mask, labels)
    lp = lambda z: F.log_softmax(z, dim=-1, dtype=torch.float32).transpose(0, 1)
    ctc = lambda z: F.ctc_loss(z, ft, il, tl, blank=cfg.pad_token_id, reduction=cfg.ctc_loss_reduction,
                               zero_infinity=cfg.ctc_zero_infinity)
    out["ctc_unmask"] = ctc(lp(logits_unmask))
    out["ctc_lm"] = ctc(lp(logits))
    out["ctc_r"] = ctc(_GRL.apply(lp(logits_r), lam))
    w = [0.1, 0.9] if args.W_LOSS is None else args.W_LOSS
    if args.AD_loss == "cel":
        lf = lambda z: F.cross_entropy(z, dementia_labels)
    else:
        wk = {"recall": w, "prec": [0.1, 0.9]}.get(args.AD_loss, [0.5, 0.5])
        lf = lambda z: recall_loss(z, dementia_labels, wk, args.AD_loss)
    out["ad"] = lf(ad_mean)
    out["ad_unmask"] = lf(ad_mean_unmask)
    out["ad_rev"] = lf(ad_mean_r)
    n = lm_masked.shape[0] * lm_masked.shape[1]
    scores = torch.cat((lm_masked.reshape(n, d), ad_masked.reshape(n, d)), 0)
    am_labels = torch.cat((torch.zeros(n, dtype=torch.long, device=scores.device), torch.ones(n, dtype=torch.long, device=scores.device)))
    out["cosface"] = cosface_loss(scores, am_labels, sd["criterion_similar.fc.weight"])
    if args.STAGE == 0:
        out["loss"] = out["ctc_unmask"]
    elif args.STAGE == 1:
        out["loss"] = out["ad_unmask"]
    else:
        out["loss"] = out["ctc_lm"] + out["ad_rev"] + out["ctc_r"] + out["ad"] + out["cosface"]
    return out


def default_args(stage: int = 2, **kw):
    a = dict(LAMBDA=0.5, TOGGLE_RATIO=0, GS_TAU=1, AD_loss="recall", W_LOSS=None, STAGE=stage)
    a.update(kw)
    return SimpleNamespace(**a)


def gen_gumbels(shape, seed: int) -> torch.Tensor:
    """-log(Exp(1)) exactly as models.py:112, from a private CPU generator."""
    g = torch.Generator().manual_seed(seed)
    return -torch.empty(shape).exponential_(generator=g).log()


def average_weights(w: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    """federated/src/utils.py:240-249: uniform mean, left-to-right fp32 sum then divide by N."""
    avg = copy.deepcopy(w[0])
    for k in avg.keys():
        for i in range(1, len(w)):
            avg[k] += w[i][k]
        avg[k] = torch.div(avg[k], len(w))
    return avg
