"""KGEModel — drop-in for KGE/models/base_model/BaseModel.py of the reference.

Same public surface (constructor arguments :29-56, `train` :58-61, `evaluate` :578,
`get_rank` :620, `score_hrt` :410, `restore_model_weights` :656, attribute
`model_weights`), but the two hot loops are library calls:

  * one `kge_train_step` per batch replaces `__run_single_batch` (:293-330): constraint,
    Philox negative sampling, gather, projection, score, loss, analytic backward,
    per-variable clip_by_norm and the Keras-semantics Adam update all run in the CUDA
    kernels of libkge_b200.so; the loss stays on the device and is read once per epoch.
  * one `kge_rank` per query block replaces the per-triple `get_rank` loop (:602-604,
    :620-654): integer "better-scoring candidate" counts with a known-positive filter.

Host control flow (epochs, early stopping, checkpoints, logging) is ordinary Python.
There is no TensorFlow and no CPU fallback: without the CUDA library every compute
entry point raises.
"""
import datetime
import logging
import math
import os

import numpy as np
import torch

from ... import metrics as _metrics
from ...data_utils import DeviceBatchStream, calculate_data_size, load_triples
from ...ns_strategy import TypedStrategy, UniformStrategy, _type_csr, check_typed_pools
from ...optimizers import adam_fields
from ..._b200 import dp as _dp, engine as _engine, lib as _lib, shard as _shard, spec as _spec

logging.getLogger().setLevel(logging.INFO)


def _now():
    return str(datetime.datetime.now())


class WeightView:
    """One entry of `model_weights`: a window on a device table that keeps the handful of
    tf.Variable conveniences callers use (`.numpy()`, `.shape`, `.assign()`)."""

    def __init__(self, name, table):
        self.name, self._table = name, table

    @property
    def tensor(self):
        return self._table.view

    @property
    def shape(self):
        return tuple(self._table.shape)

    def numpy(self):
        return self._table.view.detach().cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a.astype(dtype) if dtype is not None else a

    def assign(self, value):
        v = value.tensor if isinstance(value, WeightView) else value
        v = torch.as_tensor(np.asarray(v) if not torch.is_tensor(v) else v, dtype=torch.float32)
        if tuple(v.shape) != self.shape:
            raise ValueError(f"{self.name}: shape {tuple(v.shape)} != {self.shape}")
        if hasattr(self._table, "assign"):          # sharded table: every rank keeps its own rows
            self._table.assign(v)
            return self
        self._table.view.copy_(v.to(self._table.storage.device))
        self._table.touch()          # the training step's shadow of this table is stale now
        return self

    def __repr__(self):
        return f"WeightView({self.name}, shape={self.shape})"


class ShardedTableView:
    """The table object behind a WeightView when the entity tables are sharded over the GPUs
    (_b200/shard.py): reading gathers the full table, assigning scatters it."""

    def __init__(self, trainer, slot, shape):
        self._trainer, self._slot, self.shape = trainer, slot, tuple(shape)

    @property
    def view(self):
        return self._trainer.full_table(self._slot)

    def assign(self, value):
        self._trainer.assign_table(self._slot, value)


class _CheckpointWriter:
    """Checkpoint files written off the training thread (one GPU, unsharded tables).  `snapshot` copies a device
    tensor into a pinned host buffer in STREAM order — the snapshot is the state after the last enqueued step, and
    the steps enqueued afterwards run while it drains; a worker thread waits for the copies and writes the files
    (to a temporary name, then renamed) under the next epoch.  One write in flight: the next save, any load and
    the end of train() join it first, so the pinned buffers are never overwritten under a writer."""

    def __init__(self):
        self.thread, self.buf, self.err = None, {}, None

    def snapshot(self, key, t):
        b = self.buf.get(key)
        if b is None or b.shape != t.shape or b.dtype != t.dtype:
            b = self.buf[key] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        b.copy_(t, non_blocking=True)
        return b

    def submit(self, jobs):
        import threading
        done = torch.cuda.Event()
        done.record()

        def work():
            try:
                done.synchronize()
                for obj, path in jobs:
                    torch.save(obj, path + ".tmp")
                    os.replace(path + ".tmp", path)
            except Exception as e:          # surfaced by the next join()
                self.err = e
        self.thread = threading.Thread(target=work, name="kge-checkpoint")
        self.thread.start()

    def join(self):
        if self.thread is not None:
            self.thread.join()
            self.thread = None
        if self.err is not None:
            e, self.err = self.err, None
            raise e


class KGEModel:
    """Base of all models.  Subclasses declare `MODEL` (kernel family name) and
    `_tables(n_ent, n_rel)` -> {key: (shape, init)} with init = limit | "eye"."""

    MODEL = None
    HAS_SCORE_FN = True

    def __init__(self, embedding_params, negative_ratio, corrupt_side, loss_fn, ns_strategy, n_workers):
        assert corrupt_side in ["h+t", "h", "t"], "Invalid corrupt_side, valid options: 'h+t', 'h', 't'"
        self.embedding_params = embedding_params
        self.negative_ratio = negative_ratio
        self.corrupt_side = corrupt_side
        self.loss_fn = loss_fn
        self.ns_strategy = ns_strategy
        self._n_workers = n_workers          # kept for signature parity; sampling is a kernel
        # one process per GPU under torchrun: LOCAL_RANK picks the device
        self.device = os.environ.get("KGE_DEVICE", "cuda:%d" % int(os.environ.get("LOCAL_RANK", "0")))
        self.metadata = None
        self._model_weights_initial = None
        self._engine = None
        self._sharded = None                 # _b200.shard.ShardedTrainer when the entity tables are sharded
        self._optimizer = "Adam"
        self.seed = None

    # ---- subclass hooks ---------------------------------------------------------------
    def _dims(self):
        """(d_ent, d_rel)"""
        d = self.embedding_params.get("embedding_size")
        assert d is not None, "'embedding_size' should be given in embedding_params when using %s" % self.MODEL
        return d, d

    def _tables(self, n_ent, n_rel):
        raise NotImplementedError("subclass of KGEModel should implement _tables()")

    def _post_init(self):
        """init-time constraints (TransE.py:108-109, RESCAL.py:109-115)."""

    def _constraint_flag(self):
        return bool(getattr(self, "constraint", False))

    def _rotate_limit(self):
        return None

    # ---- weights ----------------------------------------------------------------------
    def _sizes(self):
        return len(self.metadata["ind2ent"]), len(self.metadata["ind2rel"])

    def _check_model_weights(self, model_weights):
        n_ent, n_rel = self._sizes()
        for key, (shape, _) in self._tables(n_ent, n_rel).items():
            assert model_weights.get(key) is not None, "'%s' should be given in model_weights" % key
            assert list(np.shape(_plain(model_weights[key]))) == list(shape), \
                "shape of '%s' should be %s" % (key, tuple(shape))

    def _init_embeddings(self, seed):
        """Uniform(+-limit) tables or `model_weights_initial`; builds the engine.
        Every table has its own stream of the seeded generator: the reference's quirk of
        identical same-shaped tables under a seed is not replicated."""
        n_ent, n_rel = self._sizes()
        spec = self._tables(n_ent, n_rel)
        values = {}
        if self._model_weights_initial is None:
            gen = torch.Generator()
            if seed is None:
                gen.seed()
            else:
                gen.manual_seed(int(seed))
            for key, (shape, init) in spec.items():
                if init == "eye":
                    values[key] = torch.eye(shape[1], shape[2]).repeat(shape[0], 1, 1)
                else:
                    values[key] = (torch.rand(shape, generator=gen, dtype=torch.float32) * 2 - 1) * float(init)
        else:
            self._check_model_weights(self._model_weights_initial)
            values = {k: torch.as_tensor(_plain(self._model_weights_initial[k]), dtype=torch.float32) for k in spec}
        self._build_engine(values)
        self._post_init()

    def _build_engine(self, values):
        n_ent, n_rel = self._sizes()
        d_ent, d_rel = self._dims()
        cfg = _spec.make_cfg(self.MODEL, n_ent, n_rel, d_ent, d_rel,
                             score_fn=getattr(self, "score_fn", None) if self.HAS_SCORE_FN else None,
                             loss_fn=self.loss_fn, negative_ratio=self.negative_ratio,
                             corrupt_side=self.corrupt_side, constraint=self._constraint_flag(),
                             constraint_weight=getattr(self, "constraint_weight", 1.0),
                             rotate_limit=self._rotate_limit())
        cfg.update(adam_fields(self._optimizer))
        slots = _spec.SLOTS[self.MODEL]
        self._sharded = None
        world = self._shard_world(cfg, values)
        if world:
            # entity tables (w, m, v, last) sharded over the ranks; relation tables replicated
            self._sharded = _shard.ShardedTrainer(cfg, {slots[k]: v for k, v in values.items()},
                                                  getattr(self, "batch_size", None) or 1, self.device,
                                                  emulate=world if _dp.world()[1] == 1 else 0)
            self._engine = None
            self.model_weights = {k: WeightView(k, ShardedTableView(self._sharded, slots[k], values[k].shape))
                                  for k in values}
            return
        self._engine = _engine.Engine(cfg, {slots[k]: v for k, v in values.items()}, device=self.device,
                                      lazy=self._lazy_rows(cfg, values))
        self.model_weights = {k: WeightView(k, self._engine.tab[slots[k]]) for k in values}

    def _row_sparse_eligible(self, cfg):
        """Models / settings the row-lazy and sharded steps cover: vector models without a dense regulariser
        (any of the optimizers: Adam, SGD, Adagrad)."""
        return self.MODEL in ("UM", "TransE", "TransD", "RotatE", "DistMult") or \
            (self.MODEL == "TransH" and not cfg["constraint"])

    def _shard_world(self, cfg, values):
        """Number of ranks the entity tables are sharded over, 0 = replicated tables.  Under torchrun
        (or KGE_B200_SHARD_EMULATE=G ranks emulated on one device) the tables are sharded when they exceed
        KGE_B200_SHARD_BYTES (default 512 MB: far beyond L2, the dense per-batch passes would dominate) or when
        KGE_B200_SHARD=1 forces it; KGE_B200_SHARD=0 keeps full replicas + the data-parallel step."""
        emu = int(os.environ.get("KGE_B200_SHARD_EMULATE", "0"))
        world = _dp.world()[1] if _dp.world()[1] > 1 else (emu if emu > 1 else 1)
        flag = os.environ.get("KGE_B200_SHARD", "auto")
        if world < 2 or flag == "0" or getattr(self, "batch_size", None) is None:
            return 0
        if not self._row_sparse_eligible(cfg):
            if flag == "1":
                raise NotImplementedError("KGE_B200_SHARD=1: sharded entity tables cover the vector models without a "
                                          "dense regulariser, with Adam")
            return 0
        if flag == "1":
            return world
        ent_bytes = 4 * sum(int(np.prod(v.shape)) for k, v in values.items() if k.startswith("ent"))
        return world if ent_bytes >= int(os.environ.get("KGE_B200_SHARD_BYTES", str(512 << 20))) else 0

    def _lazy_rows(self, cfg, values):
        """Row-lazy replay of the dense per-batch table passes (kge_abi.h, kge_tables.lazy_t; results are
        bit-identical to the dense step).  KGE_B200_LAZY_ROWS=1|0 forces it; by default it is used when
        the entity table is far larger than L2 (>= 512 MB) on a single GPU, for the models it covers."""
        flag = os.environ.get("KGE_B200_LAZY_ROWS", "auto")
        if _dp.world()[1] > 1 or flag == "0":
            return False
        if not self._row_sparse_eligible(cfg):
            if flag == "1":
                raise NotImplementedError("KGE_B200_LAZY_ROWS=1: row-lazy tables cover the vector models without a "
                                          "dense regulariser")
            return False
        if flag == "1":
            return True
        ent_bytes = 4 * sum(int(np.prod(v.shape)) for k, v in values.items() if k.startswith("ent"))
        return ent_bytes >= (512 << 20)

    def restore_model_weights(self, model_weights):
        self._check_model_weights(model_weights)
        if self._engine is None and self._sharded is None:
            self._model_weights_initial = model_weights
            self._init_embeddings(self.seed)
            return
        for k, view in self.model_weights.items():
            view.assign(_plain(model_weights[k]))

    def _require_engine(self):
        """The engine scoring / ranking run on: the model's own, or — sharded entity tables — one on the
        gathered tables (ShardedTrainer.eval_engine; under torchrun every rank must make the call)."""
        if self._sharded is not None:
            return self._sharded.eval_engine()
        if self._engine is None:
            raise RuntimeError("model has no weights yet: call train() or _init_embeddings() first")
        return self._engine

    def _train_engines(self):
        return self._sharded.engines if self._sharded is not None else [self._require_engine()]

    # ---- sampler ----------------------------------------------------------------------
    def _setup_sampler(self):
        n_ent = self._sizes()[0]
        ns = self.ns_strategy
        if ns is UniformStrategy:
            self.ns_strategy = ns = UniformStrategy(sample_pool=np.arange(n_ent))
        elif ns is TypedStrategy:
            labels = self.metadata["ind2type"]
            self.metadata["type2inds"] = {t: np.array([i for i, ti in enumerate(labels) if ti == t])
                                          for t in np.unique(labels)}
            self.ns_strategy = ns = TypedStrategy(pool=None, metadata={"type2inds": self.metadata["type2inds"],
                                                                       "ind2type": labels})
        if isinstance(ns, UniformStrategy):
            pool = ns.pool_array()
            for eng in self._train_engines():
                eng.set_uniform_pool(None if np.array_equal(pool, np.arange(n_ent)) else pool)
        elif isinstance(ns, TypedStrategy):
            ent_type, off, members, pos = ns.csr()
            check_typed_pools(ent_type, off, pos)
            for eng in self._train_engines():
                eng.set_typed_pools(ent_type, off, members, pos)
        else:
            raise NotImplementedError("ns_strategy must be UniformStrategy or TypedStrategy (class or instance): "
                                      "negatives are drawn inside the CUDA step kernel")

    # ---- training ---------------------------------------------------------------------
    def train(self, train_X, val_X, metadata, epochs, batch_size, early_stopping_rounds=None,
              model_weights_initial=None, restore_best_weight=True, optimizer="Adam", seed=None,
              log_path="./logs", log_projector=False, resume=False):
        """Same arguments as the reference (BaseModel.py:58-61).  `seed` additionally seeds the
        Philox negative-sampling stream (the reference leaves that stream unseeded).
        `resume=True` (extension): continue from the checkpoint in `log_path` — weights, optimizer slots,
        step counters and the position in the seeded batch stream."""
        self.metadata = metadata
        self.batch_size = batch_size
        self._model_weights_initial = model_weights_initial
        self._optimizer = optimizer
        self.seed = seed
        self.log_path = log_path
        os.makedirs(log_path, exist_ok=True)
        writers = _Writers(log_path, val_X is not None, enabled=_dp.world()[0] == 0)

        logging.info("[%s] Preparing for training..." % _now())
        n_train = calculate_data_size(train_X)
        n_batches = int(np.ceil(n_train / batch_size))
        n_val_batches = int(np.ceil(calculate_data_size(val_X) / batch_size)) if val_X is not None else 0
        train_iter = DeviceBatchStream(train_X, batch_size, shuffle=True, seed=seed, device=self.device)
        val_iter = DeviceBatchStream(val_X, batch_size, shuffle=False, device=self.device) if val_X is not None else None
        self._init_embeddings(seed=seed)
        self._setup_sampler()
        eng, sharded = self._engine, self._sharded
        device = torch.device(self.device)
        ns_seed = int(seed) if seed is not None else int.from_bytes(os.urandom(7), "little")
        # multi-GPU (torch.distributed initialised by the launcher): every rank walks the same seeded batch
        # sequence and owns a contiguous share of each batch.  Tables that fit one GPU are replicated and the
        # gradients exchanged per step (kge_train_step_dp); entity tables beyond KGE_B200_SHARD_BYTES are
        # sharded over the ranks (kge_shard_*).  Matrix models (SE / TransR / RESCAL) split the batch too, their
        # matrix gradients travelling in the all-reduced flat buffer; only RESCAL with its dense regulariser
        # (constraint=True) has no split step: its ranks repeat the whole batch (identical replicas).
        rank, world_size = _dp.world()
        split = world_size > 1 and sharded is None and not (self.MODEL == "RESCAL" and bool(eng.cfg.constraint))
        if world_size > 1:
            assert seed is not None, "multi-GPU training needs a seed: every rank must walk the same batch sequence"
        step, first_epoch = 0, 0
        self.train_loss_history, self.val_loss_history = [], []
        self.best_step, patience = None, 0
        if resume:
            state = self._load_checkpoint(with_optimizer=True)
            step, first_epoch = int(state.get("step", 0)), int(state.get("epoch", -1)) + 1
            self.train_loss_history = list(state.get("train_loss_history", []))
            self.val_loss_history = list(state.get("val_loss_history", []))
            self.best_step, patience = state.get("best_step"), int(state.get("patience", 0))
            train_iter.skip(first_epoch * n_batches)
            if val_iter is not None:
                val_iter.skip(first_epoch * min(n_batches, n_val_batches))

        # Dense tables: the step runs as a captured CUDA graph (kge_step_graph: one launch per batch instead of ~20) and
        # the NEXT batch — the stream is one batch ahead — is grouped while this one trains (kge_train_step_next);
        # results are bit-identical to plain kge_train_step calls (tests/test_gpu_graph.py).  KGE_B200_GRAPH=0: plain calls.
        graph = None
        lookahead = sharded is None and eng is not None and not eng.lazy
        if lookahead and split:
            lo, hi = _dp.batch_shard(batch_size)
            # without validation batches nothing reads w during an epoch: only wn travels between the ranks per step
            eng.dp_w_private = val_iter is None and os.environ.get("KGE_DP_W_PRIVATE", "1") != "0"
        if lookahead and os.environ.get("KGE_B200_GRAPH", "1") != "0":
            graph = eng.step_graph_dp(hi - lo, ns_seed, batch_size, lo) if split else eng.step_graph(batch_size, ns_seed)
        nxt = next(train_iter) if lookahead else None

        logging.info("[%s] Start Training..." % _now())
        for epoch in range(first_epoch, epochs):
            t_loss = torch.zeros(n_batches, dtype=torch.float32, device=device)
            v_loss = torch.zeros(max(n_val_batches, 1), dtype=torch.float32, device=device)
            for b in range(n_batches):
                if lookahead:
                    batch, nxt = nxt, next(train_iter)
                    n_step = step + 1 + (1 if (val_iter is not None and b < n_val_batches) else 0)   # the next TRAINING step
                    cur, ahead = (batch[lo:hi], nxt[lo:hi]) if split else (batch, nxt)
                    if graph is not None:
                        t_loss[b:b + 1].copy_(graph.launch(cur, step, next_triples=ahead, next_step=n_step))
                    elif split:
                        eng.train_step_dp(cur, ns_seed, step, len(batch), lo, loss_out=t_loss[b:b + 1], next_triples=ahead,
                                          next_step=n_step)
                    else:
                        eng.train_step(cur, ns_seed, step, is_train=True, loss_out=t_loss[b:b + 1], next_triples=ahead,
                                       next_step=n_step)
                else:
                    batch = next(train_iter)
                    if sharded is not None:
                        sharded.train_step(batch, ns_seed, step, loss_out=t_loss[b:b + 1])
                    elif split:
                        lo, hi = _dp.batch_shard(len(batch))
                        eng.train_step_dp(batch[lo:hi], ns_seed, step, len(batch), lo, loss_out=t_loss[b:b + 1])
                    else:
                        eng.train_step(batch, ns_seed, step, is_train=True, loss_out=t_loss[b:b + 1])
                step += 1
                if val_iter is not None and b < n_val_batches:
                    # validation batches draw negatives and apply the constraint too.  Sharded
                    # tables: the loss is computed on the gathered replica and the constraint's rewrite of w is
                    # left to the next training step, which applies the same constraint (differs by rounding of
                    # normalise(normalise(w)) only)
                    veng = sharded.eval_engine() if sharded is not None else eng
                    veng.train_step(next(val_iter), ns_seed, step, is_train=False, loss_out=v_loss[b:b + 1])
                    step += 1
            self.train_loss_history.append(float(t_loss.sum().item()) / n_batches)       # one sync per epoch
            if eng is not None:
                eng.check_peers_alive()          # a cross-GPU barrier that timed out is fatal (kge_peer_barrier)
                eng.dp_sync_weights()            # dp_w_private: w of the rows other ranks own (collective, once per epoch)
            writers.scalar("train", self.train_loss_history[-1], epoch)
            if val_iter is not None:
                self.val_loss_history.append(float(v_loss.sum().item()) / n_val_batches)
                writers.scalar("validation", self.val_loss_history[-1], epoch)
                logging.info("epoch: %i, train loss: %f, valid loss: %f"
                             % (epoch, self.train_loss_history[-1], self.val_loss_history[-1]))
            else:
                logging.info("epoch: %i, train loss: %f" % (epoch, self.train_loss_history[-1]))
            if sharded is None:
                writers.histograms(self.model_weights, epoch)
            self._progress = {"epoch": epoch, "step": step}

            if early_stopping_rounds is not None:
                assert val_X is not None, "val_X should be given if want to check early stopping."
                stop, patience = self._check_early_stopping(self.val_loss_history, patience, early_stopping_rounds,
                                                            epoch, restore_best_weight)
                if stop:
                    logging.info("[%s] Val loss does not improve within %i iterations, trigger early stopping."
                                 % (_now(), early_stopping_rounds))
                    break
            else:
                self._save_checkpoint(patience=patience)
            # rank-0-only host work (checkpoint file, histograms) must not race the in-stream peer barriers of
            # the next step: every rank waits here
            if world_size > 1:
                torch.distributed.barrier()
        self._join_checkpoint()           # the last checkpoint's files are complete before train() returns
        if log_projector:
            writers.projector(self.model_weights, self.metadata)
        writers.close()
        logging.info("[%s] Finished training!" % _now())

    fit = train      # alias used by BASELINE.json's prose (the reference method is train())

    def _check_early_stopping(self, history, patience, patience_max, step, restore_best_weight):
        if step == 0 or self.best_step is None:
            self.best_step = step
            self._save_checkpoint(patience=patience)
            return False, patience
        if history[step] >= history[self.best_step]:
            patience += 1
        else:
            patience, self.best_step = 0, step
            self._save_checkpoint(patience=patience)
        if patience == patience_max:
            if restore_best_weight:
                logging.info("[%s] Restore best weight from %i to %i step." % (_now(), step, self.best_step))
                self._load_checkpoint(with_optimizer=False)
            return True, patience
        return False, patience

    # ---- checkpoints (reference: tf.train.CheckpointManager, BaseModel.py:247-253, 515-537) ------
    def _ckpt_path(self):
        return os.path.join(self.log_path, "ckpt.pt")

    def _opt_path(self, rank):
        return os.path.join(self.log_path, "ckpt.opt.rank%d.pt" % rank)

    def _save_checkpoint(self, patience=0):
        """max_to_keep=1 like the reference's CheckpointManager (BaseModel.py:253).  `ckpt.pt` (rank 0): the
        weights — all the reference saves — plus the step counters; `ckpt.opt.rank<r>.pt` (EVERY rank): the
        optimizer slots as that rank holds them (whole tables on one GPU; the rows a rank owns under the
        owner-partitioned exchange; its shard under sharded tables) with the row-lazy stamps.  Under torchrun
        every rank must call this (reading a sharded table is a collective)."""
        rank, world = _dp.world()
        self._join_checkpoint()
        # one GPU, unsharded: pinned snapshots in stream order + a writer thread (the files are complete when
        # train() returns or the checkpoint is loaded); otherwise (collective reads) written here
        bg = world == 1 and self._sharded is None and os.environ.get("KGE_B200_ASYNC_CKPT", "1") != "0"
        if bg:        # the snapshot lives in pinned host memory (weights + two optimizer slots): bounded
            nbytes = 3 * sum(int(np.prod(v.shape)) * 4 for v in self.model_weights.values())
            bg = nbytes <= int(os.environ.get("KGE_B200_ASYNC_CKPT_MAX_MB", "1024")) << 20
        if bg and getattr(self, "_ckpt_writer", None) is None:
            self._ckpt_writer = _CheckpointWriter()
        host = (lambda key, t: self._ckpt_writer.snapshot(key, t.detach())) if bg else (lambda key, t: t.detach().cpu())
        engines = self._train_engines()
        for e in engines:
            e.flush()
        weights = {k: host("w." + k, v.tensor) for k, v in self.model_weights.items()}       # collective when sharded
        prog = getattr(self, "_progress", {"epoch": -1, "step": 0})
        jobs = []
        if rank == 0:
            jobs.append(({"weights": weights, "adam_t": engines[0].adam_t, "epoch": prog["epoch"], "step": prog["step"],
                          "world": world, "sharded": self._sharded is not None, "n_engines": len(engines),
                          "train_loss_history": list(getattr(self, "train_loss_history", [])),
                          "val_loss_history": list(getattr(self, "val_loss_history", [])),
                          "best_step": getattr(self, "best_step", None), "patience": int(patience)}, self._ckpt_path()))
        opt = []
        for i, e in enumerate(engines):
            opt.append({"adam_t": e.adam_t, "lazy_t": e.lazy_t,
                        "m": {s: host("m.%d.%d" % (i, s), t.m) for s, t in e.tab.items() if t.m is not None},
                        "v": {s: host("v.%d.%d" % (i, s), t.v) for s, t in e.tab.items() if t.v is not None},
                        "last": {s: host("l.%d.%d" % (i, s), t.last) for s, t in e.tab.items() if t.last is not None}})
        jobs.append(({"world": world, "sharded": self._sharded is not None, "engines": opt}, self._opt_path(rank)))
        if bg:
            self._ckpt_writer.submit(jobs)
        else:
            for obj, path in jobs:
                torch.save(obj, path)

    def _join_checkpoint(self):
        w = getattr(self, "_ckpt_writer", None)
        if w is not None:
            w.join()

    def _load_checkpoint(self, with_optimizer=True):
        """Weights from `ckpt.pt`; with_optimizer: also this rank's optimizer slots, Adam step and lazy stamps
        (only when the file was written by the same world size / sharding, else the slots restart at zero with
        a warning).  Returns the checkpoint's bookkeeping dict."""
        rank, world = _dp.world()
        self._join_checkpoint()
        if world > 1:
            torch.distributed.barrier()
        state = torch.load(self._ckpt_path())
        for k, v in state["weights"].items():
            self.model_weights[k].assign(v)
        if not with_optimizer:
            return state
        engines = self._train_engines()
        path = self._opt_path(rank)
        ok = os.path.exists(path)
        if ok:
            opt = torch.load(path)
            ok = opt["world"] == world and opt["sharded"] == (self._sharded is not None) and len(opt["engines"]) == len(engines)
        if not ok:
            logging.warning("checkpoint has no optimizer state for this rank / layout: slots restart at zero")
            return state
        for e, o in zip(engines, opt["engines"]):
            e.flush()
            for t in e.tab.values():
                t.ensure_slots(e.acc_init)
            for s, m in o["m"].items():
                e.tab[s].m.copy_(m.to(e.device))
                e.tab[s].v.copy_(o["v"][s].to(e.device))
            for s, l in o["last"].items():
                e.tab[s].last.copy_(l.to(e.device))
            e.adam_t, e.lazy_t = int(o["adam_t"]), int(o["lazy_t"])
        return state

    def load_checkpoint(self, log_path=None, with_optimizer=True):
        """Public restore of the checkpoint in `log_path` (default: the last train() call's) into a model whose
        tables exist (after train() or restore_model_weights())."""
        if log_path is not None:
            self.log_path = log_path
        return self._load_checkpoint(with_optimizer=with_optimizer)

    # ---- scoring / ranking ------------------------------------------------------------
    def score_hrt(self, h, r, t):
        """Scores of triples given as id vectors; `h=None` or `t=None` scores every entity on
        that side against scalar ids (BaseModel.py:410-430).  Returns a torch tensor on the device."""
        assert not (h is None and t is None), "h and t should not be None simultaneously"
        eng, n_ent = self._require_engine(), self._sizes()[0]

        def ids(x):
            return np.atleast_1d(np.asarray(_plain(x))).astype(np.int64)
        if h is None or t is None:
            r_, o_ = ids(r), ids(t if h is None else h)
            assert r_.size == 1 and o_.size == 1, "r and the given entity must be scalars when one side is None"
            every = np.arange(n_ent, dtype=np.int64)
            rep_r, rep_o = np.full(n_ent, r_[0]), np.full(n_ent, o_[0])
            return eng.score(every, rep_r, rep_o) if h is None else eng.score(rep_o, rep_r, every)
        return eng.score(ids(h), ids(r), ids(t))

    def _rank_block(self, X, corrupt_side, positive_X):
        eng, n_ent = self._require_engine(), self._sizes()[0]
        assert corrupt_side in ("h", "t"), "corrupt_side can be 'h' or 't'"
        X = np.asarray(_plain(X)).reshape(-1, 3).astype(np.int64)
        side = _lib.SIDE_H if corrupt_side == "h" else _lib.SIDE_T
        if len(X) == 0:                   # an empty split: the reference's loop (BaseModel.py:602-604) yields no ranks
            return np.zeros(0, dtype=np.int64)
        # the known-positive index is built ONCE per call (one device sort of positive_X), every block of
        # queries then costs two binary searches per query (kge_filter_lookup)
        index = None if positive_X is None else eng.filter_index(np.asarray(_plain(positive_X)), corrupt_side)
        counts = []
        # multi-GPU: each rank counts over its slice of the entity table, counts are summed
        e_lo, e_hi = _dp.entity_shard(n_ent)
        Xd = eng._ids(X)                  # one upload of the queries; the blocks are views of it
        for lo in range(0, len(X), 16384):
            blk = Xd[lo:lo + 16384]
            off = ids = None
            if index is not None:
                off, ids = index.lookup(blk)
            counts.append(_dp.allreduce_counts(eng.rank_counts(blk, side, off, ids, ent_lo=e_lo, ent_hi=e_hi)))
        # the host waits once, after the last block is enqueued
        return torch.cat(counts).cpu().numpy().astype(np.int64) + 1

    def get_rank(self, x, positive_X, corrupt_side):
        """rank = 1 + #{candidates scoring strictly higher}, known positives excluded
        (BaseModel.py:620-654).  The count is 32-bit, not the reference's int16."""
        return np.int_(self._rank_block(x, corrupt_side, positive_X)[0])

    def evaluate(self, eval_X, corrupt_side, positive_X=None):
        ranks = self._rank_block(load_triples(eval_X), corrupt_side, positive_X)
        self.last_ranks = ranks
        return _metrics.summary(ranks)


def _plain(x):
    """WeightView / torch tensor / tf-like object -> numpy-compatible."""
    if isinstance(x, WeightView):
        return x.numpy()
    if torch.is_tensor(x):
        return x.detach().cpu().numpy()
    if hasattr(x, "numpy") and not isinstance(x, np.ndarray):
        return x.numpy()
    return x


class _Writers:
    """Per-epoch scalars / histograms (BaseModel.py:444-483) and the projector export
    (:542-576) through torch.utils.tensorboard when tensorboard is importable; otherwise
    silently off — observability only, never on the hot path."""

    def __init__(self, log_path, with_val, enabled=True):
        self.log_path, self.w = log_path, {}
        if not enabled:          # only rank 0 writes event files
            return
        try:
            from torch.utils.tensorboard import SummaryWriter
            self.w["root"] = SummaryWriter(log_path)
            self.w["train"] = SummaryWriter(os.path.join(log_path, "scalar", "train"))
            if with_val:
                self.w["validation"] = SummaryWriter(os.path.join(log_path, "scalar", "validation"))
        except Exception:
            self.w = {}

    def scalar(self, which, value, step):
        if which in self.w:
            self.w[which].add_scalar("loss", value, step)

    def histograms(self, weights, step, buckets=30):
        """tf.summary.histogram's 30 equal-width buckets (BaseModel.py:463-483), counted on the device: only the
        bucket counts and four moments cross to the host (the table itself stays where it is)."""
        if "root" not in self.w:
            return
        for k, v in weights.items():
            t = v.tensor.detach()
            if t.numel() == 0 or t.numel() > (1 << 28):
                continue
            if not t.is_cuda:
                self.w["root"].add_histogram(k, v.numpy(), step)
                continue
            t = t.float().reshape(-1)
            lo, hi = torch.aminmax(t)
            stats = torch.stack([lo.double(), hi.double(), t.sum(dtype=torch.float64), (t * t).sum(dtype=torch.float64)]).cpu()
            lo_f, hi_f = float(stats[0]), float(stats[1])
            if not (math.isfinite(lo_f) and math.isfinite(hi_f)):
                continue
            if hi_f <= lo_f:
                hi_f = lo_f + 1e-12 + abs(lo_f) * 1e-6
            counts = torch.histc(t, bins=buckets, min=lo_f, max=hi_f).cpu().numpy()
            limits = np.linspace(lo_f, hi_f, buckets + 1)[1:]
            self.w["root"].add_histogram_raw(k, min=lo_f, max=hi_f, num=int(t.numel()), sum=float(stats[2]),
                                             sum_squares=float(stats[3]), bucket_limits=limits.tolist(),
                                             bucket_counts=counts.tolist(), global_step=step)

    def projector(self, weights, metadata):
        """metadata + embedding tsv files and a projector_config.pbtxt in log_path."""
        blocks = []
        for key, names in (("ent_emb", metadata.get("ind2ent")), ("rel_emb", metadata.get("ind2rel"))):
            if key not in weights or names is None:
                continue
            mat = weights[key].numpy().reshape(len(names), -1)
            np.savetxt(os.path.join(self.log_path, key + ".tsv"), mat, delimiter="\t", fmt="%.6g")
            with open(os.path.join(self.log_path, key + "_metadata.tsv"), "w") as fh:
                fh.write("\n".join(str(n) for n in names) + "\n")
            blocks.append('embeddings {\n  tensor_name: "%s"\n  tensor_path: "%s.tsv"\n  metadata_path: "%s_metadata.tsv"\n}\n'
                          % (key, key, key))
        with open(os.path.join(self.log_path, "projector_config.pbtxt"), "w") as fh:
            fh.write("".join(blocks))

    def close(self):
        for w in self.w.values():
            w.close()
