"""Look-ahead grouping (kge_train_step_next / _host_next) and the captured step graph (kge_step_graph) against
the plain kge_train_step: the results must be bit-identical — losses at every step, every table and optimizer
slot at the end — because only WHEN the grouping runs and HOW the launches are issued changes."""
import numpy as np
import pytest
import torch

from helpers import make_pair, rand_triples
from oracle import ref_ops as O

pytestmark = pytest.mark.gpu

N, R = 500, 13
CASES = [("DistMult", 48, None, "bce", 8, "h+t"), ("DistMult", 200, None, "bce", 64, "h+t"), ("DistMult", 50, None, "square", 3, "t"),
         ("TransE", 50, "l1", "hinge", 4, "h+t"), ("TransD", 40, "l2", "hinge", 2, "h"), ("RotatE", 30, "l1", "selfadv", 6, "h+t"),
         ("TransR", 24, "l2", "hinge", 2, "h+t"), ("RESCAL", 16, None, "bce", 2, "t")]
LOSSES = {"hinge": O.PairwiseHingeLoss(1.0), "bce": O.BinaryCrossEntropyLoss(), "square": O.SquareErrorLoss(),
          "selfadv": O.SelfAdversarialNegativeSamplingLoss(3.0, 1.0)}
SCORES = {"l1": O.LpDistance(1), "l2": O.LpDistance(2), None: None}


def _engine(case):
    model, d, score, loss, k, side = case
    _, _, eng = make_pair(model, N, R, d, None, score=SCORES[score], loss=LOSSES[loss], k=k, side=side)
    return eng


def _state(eng):
    out = {}
    for s, pt in eng.tab.items():
        out[("w", s)] = pt.storage.clone()
        if pt.m is not None:
            out[("m", s)], out[("v", s)] = pt.m.clone(), pt.v.clone()
    return out


def _assert_same(a, b, where):
    for key in a:
        assert torch.equal(a[key], b[key]), (where, key, float((a[key] - b[key]).abs().max()))


def _batches(n, B, seed=5):
    rng = np.random.default_rng(seed)
    return [torch.from_numpy(rand_triples(rng, B, N, R)) for _ in range(n)]


@pytest.mark.parametrize("case", CASES, ids=[c[0] + "-" + str(c[1]) for c in CASES])
def test_lookahead_equals_plain_bitwise(case):
    B, steps = 96, 6
    Xs = [x.cuda() for x in _batches(steps, B)]
    plain, ahead = _engine(case), _engine(case)
    lp, la = [], []
    for i in range(steps):
        lp.append(float(plain.train_step(Xs[i], 7, i).item()))
        nxt = Xs[i + 1] if i + 1 < steps and i != 2 else None        # step 3 groups inline: the look-ahead may be dropped any time
        la.append(float(ahead.train_step(Xs[i], 7, i, next_triples=nxt).item()))
    assert lp == la, (lp, la)
    _assert_same(_state(plain), _state(ahead), case)


def test_lookahead_not_taken_up_is_harmless():
    case = CASES[0]
    B = 64
    Xs = [x.cuda() for x in _batches(4, B)]
    plain, ahead = _engine(case), _engine(case)
    for i in range(3):
        plain.train_step(Xs[i], 7, i)
        ahead.train_step(Xs[i], 7, i, next_triples=Xs[3], next_step=99)       # announces a batch that never comes
    _assert_same(_state(plain), _state(ahead), "look-ahead ignored")


@pytest.mark.parametrize("case", CASES[:5], ids=[c[0] + "-" + str(c[1]) for c in CASES[:5]])
def test_host_lookahead_equals_plain_bitwise(case):
    B, steps = 80, 5
    Xh = [x.pin_memory() for x in _batches(steps, B)]
    plain, ahead = _engine(case), _engine(case)
    l1, l2 = torch.zeros(1).pin_memory(), torch.zeros(1).pin_memory()
    for i in range(steps):
        plain.train_step_host(Xh[i], 7, i, l1)
        ahead.train_step_host(Xh[i], 7, i, l2, next_host=Xh[i + 1] if i + 1 < steps else None)
        torch.cuda.synchronize()
        assert float(l1.item()) == float(l2.item()), (i, float(l1.item()), float(l2.item()))
    _assert_same(_state(plain), _state(ahead), case)


@pytest.mark.parametrize("lookahead", [False, True])
@pytest.mark.parametrize("case", CASES, ids=[c[0] + "-" + str(c[1]) for c in CASES])
def test_step_graph_equals_plain_bitwise(case, lookahead):
    """Captured graph vs kge_train_step over the first steps (the shadow table becomes valid after step 1, so two
    variants are captured), from host batches, with and without the look-ahead batch."""
    B, steps = 96, 7
    Xh = [x.pin_memory() for x in _batches(steps, B)]
    plain, gr = _engine(case), _engine(case)
    G = gr.step_graph(B, 7, lookahead=lookahead)
    lh = torch.zeros(1).pin_memory()
    for i in range(steps):
        want = float(plain.train_step(Xh[i].cuda(), 7, i).item())
        nxt = Xh[i + 1] if lookahead and i + 1 < steps and i != 3 else None
        G.launch(Xh[i], i, next_triples=nxt, host_loss=lh)
        torch.cuda.synchronize()
        assert float(lh.item()) == want, (i, float(lh.item()), want)
    _assert_same(_state(plain), _state(gr), case)
    # the graph and the plain call can be mixed on one engine (same tables, Adam step counter shared)
    X = Xh[0].cuda()
    a = float(plain.train_step(X, 7, 50).item())
    b = float(gr.train_step(X, 7, 50).item())
    assert a == b
    _assert_same(_state(plain), _state(gr), "mixed")


@pytest.mark.parametrize("case", [CASES[1], CASES[3], CASES[6], CASES[7]], ids=["DistMult-200", "TransE-50", "TransR-24", "RESCAL-16"])
@pytest.mark.parametrize("pinned", [True, False])
def test_step_graph_wait_loss_reads_every_step(case, pinned):
    """kge_step_graph_wait_loss: the host has each step's loss before it issues the next step, without a stream
    synchronisation.  Pinned buffers: the graph reads the look-ahead batch from host memory itself and the kernel
    that makes the loss final stores it into the host word (for RESCAL that is the dense-regulariser reduction);
    pageable buffers: the same call falls back to copies around the graph + a stream synchronisation.  Either way
    the losses and the final state are those of plain kge_train_step calls, bit for bit."""
    B, steps = 96, 9
    Xh = [x.pin_memory() if pinned else x.clone() for x in _batches(steps, B)]
    plain, gr = _engine(case), _engine(case)
    G = gr.step_graph(B, 7, lookahead=True)
    lh = torch.full((1,), float("nan"))
    lh = lh.pin_memory() if pinned else lh
    for i in range(steps):
        want = float(plain.train_step(Xh[i].cuda(), 7, i).item())
        nxt = Xh[i + 1] if i + 1 < steps and i != 4 else None              # step 5 is staged by its own launch
        G.launch(Xh[i], i, next_triples=nxt, host_loss=lh)
        G.wait_loss()
        assert float(lh[0]) == want, (i, float(lh[0]), want)
    torch.cuda.synchronize()
    _assert_same(_state(plain), _state(gr), (case, pinned))


def test_step_graph_device_batches_and_sgd():
    from KGE._b200 import lib as L
    case = ("TransE", 50, "l1", "hinge", 4, "h+t")
    B, steps = 64, 5
    Xd = [x.cuda() for x in _batches(steps, B)]
    plain, gr = _engine(case), _engine(case)
    for e in (plain, gr):
        e.cfg.optimizer = L.OPTIM_SGD if hasattr(L, "OPTIM_SGD") else 1
        e.cfg.lr = 0.05
    G = gr.step_graph(B, 3)
    for i in range(steps):
        want = float(plain.train_step(Xd[i], 3, i).item())
        got = float(G.launch(Xd[i], i, next_triples=Xd[i + 1] if i + 1 < steps else None).item())
        assert got == want, (i, got, want)
    _assert_same(_state(plain), _state(gr), "sgd")


def test_graph_and_lookahead_calls_interleaved():
    """bench.py's order: look-ahead calls, graph launches, look-ahead calls again, each leaving a grouped batch
    behind that the next kind of call does not take up."""
    case = CASES[1]
    B = 128
    Xd = [x.cuda() for x in _batches(9, B)]
    plain, mix = _engine(case), _engine(case)
    G = mix.step_graph(B, 7)
    for i in range(9):
        want = float(plain.train_step(Xd[i], 7, i).item())
        nxt = Xd[i + 1] if i + 1 < 9 else None
        if (i // 3) % 2 == 0:
            got = float(mix.train_step(Xd[i], 7, i, next_triples=nxt).item())
        else:
            got = float(G.launch(Xd[i], i, next_triples=nxt).item())
        assert got == want, (i, got, want)
    _assert_same(_state(plain), _state(mix), "interleaved")


def test_pending_lookahead_does_not_disturb_sharded_or_lazy_steps():
    """A look-ahead batch left pending by one engine (the context is shared per device) must not move the grouping
    set of steps that span several calls: sharded lookup / grad / apply, row-lazy steps (bench.py runs C2 with
    look-ahead, then C5 sharded, in one process)."""
    import test_gpu_shard as TS
    from KGE._b200 import shard as SH
    case = CASES[1]
    Xs = [x.cuda() for x in _batches(3, 64)]
    ahead = _engine(case)
    scase = TS.CASES[0]
    model, score, loss, con, d, k, side = scase
    B = 64
    batches = TS._batches(np.random.default_rng(3), 4, B)
    _, _, lazy_eng = make_pair(model, TS.N, TS.R, d, score=score, loss=loss, k=k, side=side, constraint=con, lazy=True)
    ranks = TS._sharded(scase, 1, B)
    for step, X in enumerate(batches):
        ahead.train_step(Xs[step % 3], 7, step, next_triples=Xs[(step + 1) % 3])      # leaves a grouped batch pending
        l_ref = lazy_eng.train_step(X, 5, step)
        ahead.train_step(Xs[step % 3], 7, 100 + step, next_triples=Xs[(step + 1) % 3])
        l_sh = SH.emulate_step(ranks, [X], 5, step, B, [0])
        assert float(l_ref.item()) == float(l_sh.item()), step
