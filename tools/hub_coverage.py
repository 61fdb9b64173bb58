"""How much of the gather stream a hub-column cache could serve: share of the adjacency entries that fall on the k most
referenced columns, for the k rows of B that fit (a) one SM's shared memory, (b) the distributed shared memory of a
16-CTA cluster, (c) 1/10 of the L2.  CPU only (numpy over the generated graph); DESIGN.md section 7 quotes the output.
Usage: python tools/hub_coverage.py [--shape reddit] [--F 128]  > profiles/r02_hub_coverage_<shape>.txt"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from __graft_entry__ import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="reddit")
    ap.add_argument("--F", type=int, default=128)
    a = ap.parse_args()
    mg = load_package()
    A, X, Y, C = mg.generate(a.shape, seed=1)
    cnt = np.bincount(A.colidx, minlength=A.n).astype(np.int64)
    deg = np.diff(A.rowptr.astype(np.int64))
    order = np.sort(cnt)[::-1]
    cum = np.cumsum(order)
    nnz = int(cum[-1])
    row_bytes = a.F * 4
    print(f"# {a.shape}-shape synthetic graph: n = {A.n}, entries = {nnz}, avg degree {nnz / A.n:.1f}, max row degree {int(deg.max())}, "
          f"max column count {int(order[0])}; rows of B are {row_bytes} bytes at F = {a.F}")
    print("# k most referenced columns -> share of all entries (= share of the gathers a cache holding those k rows of B would serve)")
    cases = [("shared memory of one SM (227 KB)", 227 * 1024 // row_bytes),
             ("L1 + shared memory carve-out of one SM (256 KB)", 256 * 1024 // row_bytes),
             ("DSMEM of an 8-CTA cluster (8 x 227 KB)", 8 * 227 * 1024 // row_bytes),
             ("DSMEM of a 16-CTA cluster (16 x 227 KB)", 16 * 227 * 1024 // row_bytes),
             ("1/10 of the L2 (12.6 MB)", int(12.6e6) // row_bytes)]
    for what, k in cases:
        k = min(k, A.n)
        print(f"{k:8d} columns  {100.0 * cum[k - 1] / nnz:6.2f} % of entries   {what}")
    for q in (0.5, 0.9):
        k = int(np.searchsorted(cum, q * nnz)) + 1
        print(f"{k:8d} columns carry {int(q * 100)} % of the entries ({100.0 * k / A.n:.1f} % of the columns, {k * row_bytes / 1e6:.1f} MB of B)")


if __name__ == "__main__":
    main()
