"""Indexing helpers and the batch pipeline — drop-in for KGE/data_utils.py.

`index_kg` (:23), `convert_kg_to_index` (:65), `train_test_split_no_unseen` (:102) and
`calculate_data_size` (:162) are one-off host preprocessing and stay on the host (numpy,
csv).  The tf.data iterators of `set_tf_iterator` (:176-196) are replaced by
`DeviceBatchStream`: the triples live on the GPU, each pass draws a seeded permutation
there and batches are gathered on device, so no per-step host work or H2D copy remains.
Ordering semantics are the reference's: shuffle(buffer = n) ->
repeat -> batch, i.e. every batch has exactly `batch_size` rows and batches straddle
passes; without shuffling the data wraps around.
"""
import csv
import logging
import os

import numpy as np
import torch

from .utils import check_path_exist_and_create


# ---------------------------------------------------------------------------------------
# host preprocessing
# ---------------------------------------------------------------------------------------
def _csv_files(folder):
    return [os.path.join(folder, f) for f in sorted(os.listdir(folder))]


def _read_csv(path, dtype=str):
    with open(path, newline="") as fh:
        rows = [row[:3] for row in csv.reader(fh) if row]
    return np.asarray(rows, dtype=dtype).reshape(-1, 3)


def _unique_in_order(values):
    seen = {}
    for v in values:
        if v not in seen:
            seen[v] = len(seen)
    return list(seen.keys())


def index_kg(kg_data):
    """Map entities / relations of an (n, 3) array (sorted unique order) or of a folder of
    header-less csv partitions (first-seen order) to consecutive indices."""
    if isinstance(kg_data, np.ndarray):
        entities = list(np.unique(np.append(kg_data[:, 0], kg_data[:, 2])))
        relations = list(np.unique(kg_data[:, 1]))
    else:
        ents, rels = [], []
        for f in _csv_files(kg_data):
            part = _read_csv(f)
            ents.extend(part[:, 0]); ents.extend(part[:, 2]); rels.extend(part[:, 1])
        entities, relations = _unique_in_order(ents), _unique_in_order(rels)
    return {"ent2ind": {e: i for i, e in enumerate(entities)}, "ind2ent": list(entities),
            "rel2ind": {r: i for i, r in enumerate(relations)}, "ind2rel": list(relations)}


def convert_kg_to_index(kg_data, ent2ind, rel2ind):
    """Array in -> indexed array out; folder in -> `<folder>_indexed/` with the same files."""
    def conv(part):
        cols = [[ent2ind.get(x) for x in part[:, 0]], [rel2ind.get(x) for x in part[:, 1]],
                [ent2ind.get(x) for x in part[:, 2]]]
        return np.array(cols).T
    if isinstance(kg_data, np.ndarray):
        return conv(kg_data)
    out_dir = kg_data + "_indexed"
    check_path_exist_and_create(out_dir)
    for f in _csv_files(kg_data):
        with open(os.path.join(out_dir, os.path.basename(f)), "w", newline="") as fh:
            csv.writer(fh).writerows(conv(_read_csv(f)).tolist())
    logging.info("indexed_kg has been save to %s", out_dir)


def train_test_split_no_unseen(X, test_size, seed):
    """Random split in which every entity / relation of the test part also occurs in train."""
    if isinstance(test_size, float):
        test_size = int(len(X) * test_size)
    ent, ent_cnt = np.unique(np.append(X[:, 0], X[:, 2]), return_counts=True)
    rel, rel_cnt = np.unique(X[:, 1], return_counts=True)
    left_e, left_r = dict(zip(ent, ent_cnt)), dict(zip(rel, rel_cnt))
    rnd = np.random.RandomState(seed)
    chosen = set()
    candidates = np.arange(len(X))
    loops, max_loops = 0, len(X) * 10
    while len(chosen) < test_size:
        # the reference's loop (data_utils.py:139-154) verbatim in behaviour: an index drawn a second time is tested and
        # its counters are decremented AGAIN (np.unique then collapses it), so with the same seed the split is the
        # reference's split
        i = int(rnd.choice(candidates))
        h, r, t = X[i]
        if left_e[h] > 1 and left_r[r] > 1 and left_e[t] > 1:
            left_e[h] -= 1; left_r[r] -= 1; left_e[t] -= 1
            chosen.add(i)
        loops += 1
        if loops == max_loops:
            logging.error("Cannot split a test set with desired size, please reduce the test size")
            return None
    test_id = np.array(sorted(chosen), dtype=int)
    return X[np.setdiff1d(candidates, test_id)], X[test_id]


def load_triples(data):
    """ndarray -> int64 ndarray; folder of indexed csv partitions -> concatenated int32 ndarray
    (the dtypes the reference's iterators yield, tests/test_data_utils.py:82,90)."""
    if isinstance(data, str):
        parts = [_read_csv(f, dtype=np.int64) for f in _csv_files(data)]
        return np.concatenate(parts).astype(np.int32)
    if torch.is_tensor(data):
        data = data.detach().cpu().numpy()
    return np.ascontiguousarray(np.asarray(data))


def calculate_data_size(X):
    if isinstance(X, str):
        return sum(len(_read_csv(f)) for f in _csv_files(X))
    return len(X)


# ---------------------------------------------------------------------------------------
# device batch pipeline
# ---------------------------------------------------------------------------------------
class DeviceBatchStream:
    """Infinite iterator of (batch_size, 3) id tensors resident on `device` — the reference's
    from_tensor_slices / csv -> shuffle -> repeat -> batch pipeline (data_utils.py:176-196) without the per-step
    host -> device copy.

    On a CUDA device the work is the library's (csrc/batch_stream.cu): `kge_epoch_perm` draws the order of a pass
    (stable argsort of Philox keys of (seed, pass): stateless, so `skip` / resume regenerate it) and
    `kge_batch_gather` cuts one batch; a batch that straddles two passes is two gathers.  `device="cpu"` keeps a
    torch-only iterator for host-side tooling and the CPU tests of this class (no kernels involved)."""

    def __init__(self, data, batch_size, shuffle, seed=None, device="cuda:0"):
        arr = load_triples(data)
        self.dtype = torch.int32 if arr.dtype == np.int32 else torch.int64
        self.device = torch.device(device)
        self.store = torch.as_tensor(arr.astype(np.int64)).to(self.device)        # kernels take int64 ids
        self.n, self.batch_size, self.shuffle = len(arr), int(batch_size), bool(shuffle)
        self.cuda = self.device.type == "cuda"
        self.gen = None
        self.seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (1 << 63))
        self.pass_no = 0
        if self.cuda:
            import ctypes as C
            from ._b200 import lib as L, engine as E
            self._L, self._C, self._E = L, C, E
            self._ctx = E._ctx(self.device.index if self.device.index is not None else torch.cuda.current_device())
            if self.shuffle:
                nb = C.c_size_t()
                L.check(L.load().kge_epoch_perm_bytes(self.n, C.byref(nb)))
                self._ws = torch.empty(nb.value, dtype=torch.uint8, device=self.device)
                self._perm = [torch.empty(self.n, dtype=torch.int32, device=self.device) for _ in range(2)]
        elif self.shuffle:
            self.gen = torch.Generator(device=self.device)
            self.gen.manual_seed(self.seed)
        self.order = self._new_pass()
        self.cursor = 0

    def _new_pass(self):
        k = self.pass_no
        self.pass_no += 1
        if not self.shuffle:
            return None if self.cuda else torch.arange(self.n, device=self.device)
        if not self.cuda:
            return torch.randperm(self.n, generator=self.gen, device=self.device)
        out = self._perm[k & 1]          # two buffers: a straddling batch still reads the previous pass
        with torch.cuda.device(self.device):
            self._L.check(self._L.load().kge_epoch_perm(self._ctx, self.n, self.seed, k, self._E._ptr(out), self._E._ptr(self._ws),
                                                        self._ws.numel(), self._E._stream()))
        return out

    def __iter__(self):
        return self

    def skip(self, n_batches):
        """Advance by n_batches without materialising them (resume).  CUDA: passes are a function of (seed, pass), so
        only the pass the cursor lands in is generated; CPU: the same generator draws as iterating."""
        need = int(n_batches) * self.batch_size
        if self.cuda:
            total = self.cursor + need
            passes, self.cursor = divmod(total, self.n)
            if passes > 0:
                self.pass_no += passes - 1
                self.order = self._new_pass()
            return
        while need > 0:
            take = min(need, self.n - self.cursor)
            self.cursor += take
            need -= take
            if self.cursor == self.n:
                self.order, self.cursor = self._new_pass(), 0

    def __next__(self):
        if self.cuda:
            out = torch.empty((self.batch_size, 3), dtype=torch.int64, device=self.device)
        need, pieces, done = self.batch_size, [], 0
        while need > 0:
            take = min(need, self.n - self.cursor)
            if self.cuda:       # gathers are enqueued at once: a later pass may then reuse the permutation buffer (stream order)
                with torch.cuda.device(self.device):
                    self._L.check(self._L.load().kge_batch_gather(
                        self._ctx, 1, self._E._ptr(self.store), self.n, self._E._ptr(self.order), self.cursor, take, 1,
                        self._C.c_void_p(out.data_ptr() + 24 * done), self._E._stream()))
                done += take
            else:
                pieces.append(self.order[self.cursor:self.cursor + take])
            self.cursor += take
            need -= take
            if self.cursor == self.n:
                self.order, self.cursor = self._new_pass(), 0
        if self.cuda:
            return out
        return self.store.index_select(0, pieces[0] if len(pieces) == 1 else torch.cat(pieces))


def set_tf_iterator(data, batch_size, shuffle, buffer_size=None, seed=None, device="cuda:0"):
    """Name kept for drop-in compatibility; returns a DeviceBatchStream (no TensorFlow)."""
    if shuffle:
        assert buffer_size is not None, "buffer_size must be given when shuffle is True"
    return DeviceBatchStream(data, batch_size, shuffle, seed, device)
