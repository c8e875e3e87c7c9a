"""The nine reference models — torch-CPU fp32 restatement (TEST INFRASTRUCTURE).

Weights are a dict name -> torch tensor (requires_grad set by the step).  Every
``score_hrt`` follows the reference op order; ids are int64 tensors of shape (n,).
"""
import math
import torch
from . import ref_ops as O


class Weights:
    """name -> fp32 table, plus the tape bookkeeping that reproduces TF's gradient
    *types* (SURVEY.md App. B.1): every ``lookup`` is one IndexedSlices occurrence
    list (values = d loss / d gathered rows), every ``dense`` use yields a dense
    gradient for that table."""

    def __init__(self, data=None):
        self.data = {k: v.clone().float().contiguous() for k, v in (data or {}).items()}
        self.tape = None

    def value(self, k):
        return self.data[k]

    def assign(self, k, v):
        self.data[k] = v.detach().clone()

    def start_tape(self):
        self.tape = {"sparse": [], "dense": {}}

    def stop_tape(self):
        t, self.tape = self.tape, None
        return t

    def lookup(self, k, idx):
        rows = self.data[k][idx]
        if self.tape is not None:
            rows = rows.detach().requires_grad_(True)
            self.tape["sparse"].append((k, idx, rows))
        return rows

    def dense(self, k):
        if self.tape is None:
            return self.data[k]
        if k not in self.tape["dense"]:
            self.tape["dense"][k] = self.data[k].detach().requires_grad_(True)
        return self.tape["dense"][k]

    def clone(self):
        return Weights(self.data)


class RefModel:
    """Base: constructor arguments as KGE/models/base_model/BaseModel.py:29-56."""
    name = None
    keys = ()

    def __init__(self, n_ent, n_rel, d_ent, d_rel=None, score_fn=None, loss_fn=None,
                 constraint=True, constraint_weight=1.0):
        self.n_ent, self.n_rel = n_ent, n_rel
        self.d_ent, self.d_rel = d_ent, (d_rel if d_rel is not None else d_ent)
        self.score_fn, self.loss_fn = score_fn, loss_fn
        self.constraint, self.constraint_weight = constraint, constraint_weight
        self.limit = None

    # shapes of model_weights
    def shapes(self):
        raise NotImplementedError

    def init_limits(self):
        raise NotImplementedError

    def init_weights(self, seed):
        """Uniform +-limit per table from independent seeded streams (the
        reference's identical-tables-under-seed quirk is deliberately not replicated)."""
        g = torch.Generator().manual_seed(seed)
        W = Weights()
        lim = self.init_limits()
        for k, shp in self.shapes().items():
            if lim[k] is None:           # TransR.py:125-128 rectangular identity
                W.data[k] = torch.eye(shp[1], shp[2]).repeat(shp[0], 1, 1).contiguous()
            else:
                W.data[k] = (torch.rand(shp, generator=g, dtype=torch.float32) * 2 - 1) * lim[k]
        self.post_init(W)
        return W

    def post_init(self, W):
        pass

    def constraint_loss(self, W, X):
        return 0.0


class UM(RefModel):
    """UM.py:110-137, 139-155."""
    name = "UM"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent)}

    def init_limits(self):
        return {"ent_emb": math.sqrt(6.0 / self.d_ent)}

    def score_hrt(self, W, h, r, t):
        return self.score_fn(W.lookup("ent_emb", h), W.lookup("ent_emb", t))

    def constraint_loss(self, W, X):
        if self.constraint:
            W.assign("ent_emb", O.normalized_embeddings(W.value("ent_emb"), 2, 1, -1))
        return 0.0


class TransE(RefModel):
    """TransE.py:127-155 (score), 157-173 (constraint), 108-109 (rel normalise at init)."""
    name = "TransE"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_emb": (self.n_rel, self.d_ent)}

    def init_limits(self):
        lim = 6.0 / math.sqrt(self.d_ent)
        return {"ent_emb": lim, "rel_emb": lim}

    def post_init(self, W):
        if self.constraint:
            W.assign("rel_emb", O.normalized_embeddings(W.value("rel_emb"), 2, 1, 1))

    def score_hrt(self, W, h, r, t):
        return self.score_fn(W.lookup("ent_emb", h) + W.lookup("rel_emb", r),
                             W.lookup("ent_emb", t))

    def constraint_loss(self, W, X):
        if self.constraint:
            W.assign("ent_emb", O.normalized_embeddings(W.value("ent_emb"), 2, 1, 1))
        return 0.0


class TransH(RefModel):
    """TransH.py:149-185 (score), 187-213 (constraint: dense regulariser)."""
    name = "TransH"

    def shapes(self):
        s = (self.n_rel, self.d_ent)
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_emb": s, "rel_hyper": s}

    def init_limits(self):
        lim = math.sqrt(6.0 / self.d_ent)
        return {"ent_emb": lim, "rel_emb": lim, "rel_hyper": lim}

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h), W.lookup("ent_emb", t)
        re, w = W.lookup("rel_emb", r), W.lookup("rel_hyper", r)
        hp = he - torch.sum(w * he, dim=-1, keepdim=True) * w
        tp = te - torch.sum(w * te, dim=-1, keepdim=True) * w
        return self.score_fn(hp + re, tp)

    def constraint_loss(self, W, X):
        if not self.constraint:
            return 0.0
        W.assign("rel_hyper", O.normalized_embeddings(W.value("rel_hyper"), 2, 1, 1))
        scale = O.soft_constraint(W.dense("ent_emb"), 2, 1, -1)
        rh, re_ = W.dense("rel_hyper"), W.dense("rel_emb")
        orth = torch.sum(rh * re_, dim=-1)
        # tf.norm(rel_emb, axis=-1) = sqrt(sum(x*x))
        orth = torch.pow(orth / torch.sqrt(torch.sum(re_ * re_, dim=-1)), 2) - 1e-18
        orth = torch.sum(O.clip_by_value(orth, 0.0, math.inf))
        return self.constraint_weight * (scale + orth)


class TransR(RefModel):
    """TransR.py:154-191 (score), 193-211 (constraint)."""
    name = "TransR"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_emb": (self.n_rel, self.d_rel),
                "rel_proj": (self.n_rel, self.d_ent, self.d_rel)}

    def init_limits(self):
        return {"ent_emb": math.sqrt(6.0 / self.d_ent), "rel_emb": math.sqrt(6.0 / self.d_rel),
                "rel_proj": None}

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h), W.lookup("ent_emb", t)
        re, M = W.lookup("rel_emb", r), W.lookup("rel_proj", r)
        hp = torch.matmul(he.unsqueeze(-2), M).squeeze(-2)
        tp = torch.matmul(te.unsqueeze(-2), M).squeeze(-2)
        if self.constraint:
            hp = O.clip_constraint(hp, 2, 1, -1)
            tp = O.clip_constraint(tp, 2, 1, -1)
        return self.score_fn(hp + re, tp)

    def constraint_loss(self, W, X):
        if self.constraint:
            W.assign("ent_emb", O.clip_constraint(W.value("ent_emb"), 2, 1, -1))
            W.assign("rel_emb", O.clip_constraint(W.value("rel_emb"), 2, 1, -1))
        return 0.0


class TransD(RefModel):
    """TransD.py:170-222 (score: materialised (r_p e_p^T + I) e), 224-241 (constraint)."""
    name = "TransD"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_emb": (self.n_rel, self.d_rel),
                "ent_proj": (self.n_ent, self.d_ent), "rel_proj": (self.n_rel, self.d_rel)}

    def init_limits(self):
        le, lr = math.sqrt(6.0 / self.d_ent), math.sqrt(6.0 / self.d_rel)
        return {"ent_emb": le, "ent_proj": le, "rel_emb": lr, "rel_proj": lr}

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h).unsqueeze(-1), W.lookup("ent_emb", t).unsqueeze(-1)
        hp, tp = W.lookup("ent_proj", h).unsqueeze(-1), W.lookup("ent_proj", t).unsqueeze(-1)
        re, rp = W.lookup("rel_emb", r), W.lookup("rel_proj", r).unsqueeze(-1)
        eye = torch.eye(self.d_rel, self.d_ent)
        Mh = torch.matmul(rp, hp.transpose(-1, -2)) + eye
        Mt = torch.matmul(rp, tp.transpose(-1, -2)) + eye
        hpj = torch.matmul(Mh, he).squeeze(-1)
        tpj = torch.matmul(Mt, te).squeeze(-1)
        if self.constraint:
            hpj = O.clip_constraint(hpj, 2, 1, -1)
            tpj = O.clip_constraint(tpj, 2, 1, -1)
        return self.score_fn(hpj + re, tpj)

    constraint_loss = TransR.constraint_loss


class RotatE(RefModel):
    """RotatE.py:126-165; ent_emb is (N, d, 2) re/im interleaved (RotatE.py:96)."""
    name = "RotatE"

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        margin = getattr(self.loss_fn, "margin", 6.0)      # RotatE.py:88-92
        self.limit = (margin + 2.0) / self.d_ent

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent, 2), "rel_emb": (self.n_rel, self.d_ent)}

    def init_limits(self):
        return {"ent_emb": self.limit, "rel_emb": self.limit}

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h), W.lookup("ent_emb", t)
        ph = W.lookup("rel_emb", r) / self.limit * math.pi
        c, s = torch.cos(ph), torch.sin(ph)
        hre, him = he[..., 0], he[..., 1]
        x = (hre * c - him * s, hre * s + him * c)           # complex multiply
        return self.score_fn(x, (te[..., 0], te[..., 1]))


class SE(RefModel):
    """SE.py:131-161 (score), 163-179 (constraint)."""
    name = "SE"

    def shapes(self):
        m = (self.n_rel, self.d_ent, self.d_ent)
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_proj_h": m, "rel_proj_t": m}

    def init_limits(self):
        lm = math.sqrt(3.0 / self.d_ent)
        return {"ent_emb": math.sqrt(6.0 / self.d_ent), "rel_proj_h": lm, "rel_proj_t": lm}

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h).unsqueeze(-1), W.lookup("ent_emb", t).unsqueeze(-1)
        Rh, Rt = W.lookup("rel_proj_h", r), W.lookup("rel_proj_t", r)
        return self.score_fn(torch.matmul(Rh, he).squeeze(-1), torch.matmul(Rt, te).squeeze(-1))

    constraint_loss = TransE.constraint_loss


class DistMult(RefModel):
    """DistMult.py:118-146 (score), 148-166 (constraint: sparse regulariser)."""
    name = "DistMult"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent), "rel_inter": (self.n_rel, self.d_ent)}

    def init_limits(self):
        lim = math.sqrt(6.0 / self.d_ent)
        return {"ent_emb": lim, "rel_inter": lim}

    def score_hrt(self, W, h, r, t):
        return torch.sum(W.lookup("ent_emb", h) * W.lookup("rel_inter", r)
                         * W.lookup("ent_emb", t), dim=-1)

    def constraint_loss(self, W, X):
        if not self.constraint:
            return 0.0
        W.assign("ent_emb", O.normalized_embeddings(W.value("ent_emb"), 2, 1, 1))
        ri = W.lookup("rel_inter", X[:, 1])
        return self.constraint_weight * torch.mean(O.Lp_regularization(ri, 2, -1))


class RESCAL(RefModel):
    """RESCAL.py:140-174 (score), 176-203 (constraint: dense regulariser), 109-115 (init)."""
    name = "RESCAL"

    def shapes(self):
        return {"ent_emb": (self.n_ent, self.d_ent),
                "rel_inter": (self.n_rel, self.d_ent, self.d_ent)}

    def init_limits(self):
        return {"ent_emb": math.sqrt(6.0 / self.d_ent), "rel_inter": math.sqrt(3.0 / self.d_ent)}

    def post_init(self, W):
        if self.constraint:
            W.assign("ent_emb", O.normalized_embeddings(W.value("ent_emb"), 2, 1, -1))
            W.assign("rel_inter", O.normalized_embeddings(W.value("rel_inter"), 2, 1, (1, 2)))

    def score_hrt(self, W, h, r, t):
        he, te = W.lookup("ent_emb", h).unsqueeze(-1), W.lookup("ent_emb", t).unsqueeze(-1)
        R = W.lookup("rel_inter", r)
        return torch.matmul(torch.matmul(he.transpose(-1, -2), R), te).squeeze(-1).squeeze(-1)

    def constraint_loss(self, W, X):
        if not self.constraint:
            return 0.0
        e = torch.mean(O.Lp_regularization(W.dense("ent_emb"), 2, -1))
        r = torch.mean(O.Lp_regularization(W.dense("rel_inter"), 2, (1, 2)))
        return self.constraint_weight * (e + r)


MODELS = {c.name: c for c in (UM, SE, TransE, TransH, TransR, TransD, RotatE, DistMult, RESCAL)}
