"""Diagnostic: engine vs oracle gradients at epoch 9 (weights synced), tensor-core GEMM on/off."""
import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package
mg = load_package(); mg.lib()
from oracle import oracle as orc
import torch
torch.cuda.set_device(0)
for tc in (1, 0):
    mg.lib().mggcn_tune(b"gemm_tc", tc)
    A, X, Y, C = mg.generate("oracle4k", seed=1)
    sizes = [X.shape[1], 64, 64, 64, C]
    eng, A_T = mg.build_rank_engine(A, sizes)
    ref = orc.Gcn(sizes, A.n, (A_T.rowptr, A_T.colidx, A_T.vals), (A.rowptr, A.colidx, A.vals), P=1)
    eng.set_inputs(X, Y)
    L = 4
    for ep in range(10):
        if ep == 9:
            for l in range(L):
                ref.set_tensor(f"W{l}", eng.get(f"W{l}")); ref.set_tensor(f"b{l}", eng.get(f"b{l}"))
        eng.train_forward(); ref.train_forward(X, Y)
        if ep in (0, 9):
            for l in range(L - 1):
                a, b = eng.get(f"H{l}").astype(np.float64), ref.tensor(f"H{l}").astype(np.float64).ravel()
                flips = np.nonzero((a > 0) != (b > 0))[0]
                print(f"tc={tc} ep={ep} H{l} maxdiff {np.abs(a-b).max():.3e} max {np.abs(b).max():.3e} sign flips {len(flips)} "
                      f"|ref| at flips {np.abs(b[flips])[:5]} near-zero(|Z|<2e-6) {int(((b>0)&(b<2e-6)).sum()+((b<=0)&(b>-2e-8)).sum())}")
        eng.backward(); ref.backward()
        if ep in (0, 9):
            for l in range(L):
                for nm in (f"GW{l}", f"Gb{l}"):
                    a, b = eng.get(nm).astype(np.float64), ref.tensor(nm).astype(np.float64).ravel()
                    print(f"tc={tc} ep={ep} {nm} maxdiff {np.abs(a-b).max():.3e} max {np.abs(b).max():.3e}")
        eng.adam_update(); ref.adam_update()
    eng.close()
