"""Throughput of KGEModel.train() itself on the C2 shape (DistMult d=200, BCE, 64 negatives 'h+t', B=1024, FB15k-sized
synthetic graph): whole epochs through the public API — device batch stream, step graph, per-epoch loss read,
histograms and checkpoint — timed by wall clock per epoch from the log."""
import logging, os, sys, time, tempfile
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "knowledge-graph-embedding_b200")]
from KGE import loss
from KGE.models.semantic_based.DistMult import DistMult

N, R, n_train, B, k, d = 14951, 1345, 483142, 1024, 64, 200
rng = np.random.default_rng(0)
train = np.stack([rng.integers(0, N, n_train), rng.integers(0, R, n_train), rng.integers(0, N, n_train)], 1).astype(np.int64)
meta = {"ind2ent": [str(i) for i in range(N)], "ind2rel": [str(i) for i in range(R)]}
stamps = []


class H(logging.Handler):
    def emit(self, rec):
        if "epoch:" in rec.getMessage() or "Start Training" in rec.getMessage():
            torch.cuda.synchronize()
            stamps.append(time.perf_counter())


from KGE.models.base_model import BaseModel as _BM
spent = {}


def timed(cls, name):
    fn = getattr(cls, name)

    def wrap(*a, **kw):
        torch.cuda.synchronize()
        t = time.perf_counter()
        r = fn(*a, **kw)
        torch.cuda.synchronize()
        spent.setdefault(name, []).append(time.perf_counter() - t)
        return r
    setattr(cls, name, wrap)


timed(_BM.KGEModel, "_save_checkpoint")
timed(_BM._Writers, "histograms")
timed(_BM._Writers, "scalar")
logging.getLogger().addHandler(H())
logging.getLogger().setLevel(logging.INFO)
epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 4
m = DistMult(embedding_params={"embedding_size": d}, negative_ratio=k, corrupt_side="h+t", loss_fn=loss.BinaryCrossEntropyLoss())
with tempfile.TemporaryDirectory() as tmp:
    t0 = time.perf_counter()
    m.train(train_X=train, val_X=None, metadata=meta, epochs=epochs, batch_size=B, seed=1, log_path=tmp)
    t1 = time.perf_counter()
per_epoch = np.diff(stamps)
triples = n_train * (1 + k)
print("epoch wall s:", [round(float(x), 4) for x in per_epoch])
print("triples/s per epoch (loss read + log included, checkpoint / histograms fall into the next epoch's interval): ",
      [round(triples / float(x) / 1e6, 1) for x in per_epoch], "M")
print("whole train() call: %.3f s for %d epochs = %.1f M triples/s" % (t1 - t0, epochs, epochs * triples / (t1 - t0) / 1e6))
print("per-epoch host work, ms:", {k2: [round(x * 1e3, 2) for x in v] for k2, v in spent.items()})
print("loss history", [round(x, 4) for x in m.train_loss_history])
