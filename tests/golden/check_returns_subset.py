"""Re-runs a random subset of the episodes of ``returns_<tag>.npz`` with the CURRENT oracle and compares the returns bit for bit.

    python tests/golden/check_returns_subset.py <tag> <n_episodes> <n_procs>

Used after an oracle change that cannot affect walking episodes (e.g. the left/right hip-motor contact, reachable only with both
rear hips at their limits): identical returns on the subset show the 4000-12000-episode fixtures need no regeneration."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests', 'golden'))
import numpy as np, multiprocessing as mp
import make_returns as mr
tag=sys.argv[1]; n=int(sys.argv[2]); procs=int(sys.argv[3])
fx=np.load(os.path.join(mr.OUT,f"returns_{tag}.npz"))
seeds=fx['seed'].astype(int)
rng=np.random.default_rng(0); pick=np.sort(rng.choice(len(seeds),n,replace=False))
chunks=[(tag,list(map(int,seeds[c])),int(fx['max_steps'])) for c in np.array_split(pick,procs*2)]
with mp.Pool(procs) as pool: res=[r for part in pool.map(mr._worker,chunks) for r in part]
res=sorted(res); a=np.array(res)
m={int(s):i for i,s in enumerate(seeds)}
same=sum(1 for r in res if fx['ret'][m[int(r[0])]]==r[1])
print(tag,'episodes',len(res),'bitwise identical returns',same, 'max abs diff', max(abs(fx['ret'][m[int(r[0])]]-r[1]) for r in res))
