import sys, os
sys.path[:0] = ['/root/repo', '/root/repo/knowledge-graph-embedding_b200']
import numpy as np, torch
import bench
W = bench.WORKLOAD
dev = torch.device('cuda', 0)
eng = bench.make_engine(dev)
train = bench.zipf_triples(W['n_train'], W['n_ent'], W['n_rel'], 12345)
rng = np.random.default_rng(100)
for step in range(600):
    eng.train_step(torch.from_numpy(train[rng.integers(0, len(train), 1024)]).to(dev), 12345, step)
E = eng.tab[0].view.double(); Rl = eng.tab[2].view.double()
q_all = train[np.random.default_rng(777).integers(0, len(train), 2048)]
h, r, t = [torch.from_numpy(q_all[:, i]).to(dev) for i in range(3)]
for side in ('h', 't'):
    keep = E[t] if side == 'h' else E[h]
    true = E[h] if side == 'h' else E[t]
    q = Rl[r] * keep
    pos = (q * true).sum(1)
    S = q @ E.T
    nq = q.norm(dim=1); ne = E.norm(dim=1)
    Ef = eng.tab[0].view
    hi = (Ef.view(torch.int32) + 0x1000) & ~0x1FFF      # rna-ish
    elo = (Ef - hi.view(torch.float32)).double().norm(dim=1)
    for name, band in (('eps only 8.4e-5', 8.4e-5 * nq[:, None] * ne.max()), ('2-pass', 8.4e-5 * nq[:, None] * ne.max() + nq[:, None] * elo[None, :])):
        unc = ((S - pos[:, None]).abs() <= band).sum(1).float()
        print(side, name, 'mean', unc.mean().item(), 'median', unc.median().item(), 'p99', unc.kthvalue(int(0.99 * len(unc))).values.item(), 'max', unc.max().item(), 'frac>128', (unc > 128).float().mean().item())
    print('score std / (nq*ne)', (S.std(dim=1) / (nq * ne.mean())).mean().item(), 'elo/ne', (elo / ne).mean().item(), 'ne max', ne.max().item(), 'ne mean', ne.mean().item())
