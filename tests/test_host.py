"""Host pre-processing, file formats and generator of libmggcn.so against the oracle (bit-exact) and
against the reference's fixtures (byte-exact)."""
import filecmp
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def random_csr(rng, n, m, density, weights=True):
    mask = rng.random((n, m)) < density
    rp = np.zeros(n + 1, np.uint64)
    rp[1:] = np.cumsum(mask.sum(1))
    ci = np.nonzero(mask)[1].astype(np.uint32)
    va = (rng.random(len(ci)).astype(np.float32) + 0.5) if weights else np.ones(len(ci), np.float32)
    return rp, ci, va


@pytest.mark.parametrize("toy", ["toyA", "toyB"])
def test_fixture_round_trip_is_byte_exact(mg, toy, tmp_path):
    A, info = mg.read_csr(os.path.join(GOLD, toy, "graph.bin"))
    assert (info.label_bytes, info.ordinal_bytes) == (4, 4)
    out = tmp_path / "graph.bin"
    mg.write_csr(str(out), A)
    assert filecmp.cmp(out, os.path.join(GOLD, toy, "graph.bin"), shallow=False)
    for f, dt in (("features.bin", np.float32), ("labels.bin", np.uint32), ("sets.bin", np.uint32)):
        a = mg.read_dense(os.path.join(GOLD, toy, f), dt)
        mg.write_dense(str(tmp_path / f), a)
        assert filecmp.cmp(tmp_path / f, os.path.join(GOLD, toy, f), shallow=False)
    if toy == "toyA":
        assert A.rowptr.tolist() == [0, 2, 4, 6, 8] and A.colidx.tolist() == [1, 3, 0, 2, 1, 3, 0, 2]
        assert mg.read_dense(os.path.join(GOLD, toy, "features.bin")).tolist() == [[0, 1], [1, 0], [0, 1], [1, 0]]


def test_int64_ordinal_file(mg, tmp_path):
    rng = np.random.default_rng(3)
    rp, ci, va = random_csr(rng, 50, 50, 0.1)
    A = mg.Csr(rp, ci, va, 50)
    mg.write_csr(str(tmp_path / "g64.bin"), A, ordinal_bytes=8)
    B, info = mg.read_csr(str(tmp_path / "g64.bin"))
    assert info.ordinal_bytes == 8 and info.nnz == A.nnz
    assert np.array_equal(B.rowptr, A.rowptr) and np.array_equal(B.colidx, A.colidx) and np.array_equal(B.vals, A.vals)


def test_bad_file_is_rejected(mg, tmp_path):
    p = tmp_path / "x.bin"
    p.write_bytes(b"not a graph")
    with pytest.raises(mg.MggcnError):
        mg.read_csr(str(p))
    with pytest.raises(mg.MggcnError):
        mg.read_csr(str(tmp_path / "missing.bin"))


@pytest.mark.parametrize("n,m,density,weights", [(1, 1, 1.0, True), (37, 37, 0.2, True), (200, 200, 0.05, False),
                                                 (64, 90, 0.1, True), (300, 300, 0.0, True)])
def test_normalize_transpose_bit_exact(mg, orc, n, m, density, weights):
    rng = np.random.default_rng(n * 7 + m)
    rp, ci, va = random_csr(rng, n, m, density, weights)
    for axis in (True, False):
        if density == 0.0 and not axis:
            continue
        A = mg.Csr(rp, ci, va.copy(), m)
        A.normalize(by_column=axis)
        want = orc.normalize(rp, ci, va, m, axis=axis)
        assert np.array_equal(A.vals.view(np.uint32), want.view(np.uint32))
    A = mg.Csr(rp, ci, va, m)
    T = A.transpose()
    trp, tci, tva = orc.transpose(rp, ci, va, m)
    assert np.array_equal(T.rowptr, trp) and np.array_equal(T.colidx, tci)
    assert np.array_equal(T.vals.view(np.uint32), tva.view(np.uint32))
    assert T.n == m and T.m == n


@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_block_split_bit_exact(mg, orc, P):
    rng = np.random.default_rng(P)
    n = 203  # not a multiple of P: CSR cuts are floor(i*N/P), src/main.cpp:139-141
    rp, ci, va = random_csr(rng, n, n, 0.08)
    A = mg.Csr(rp, ci, va, n)
    cuts = mg.row_cuts(n, P)
    assert cuts.tolist() == [i * n // P for i in range(P + 1)]
    blocks = A.block_rows(cuts, cuts)
    total = 0
    for i in range(P):
        only = A.block_rows(cuts, cuts, only_row=i)
        for j in range(P):
            brp, bci, bva = orc.block(rp, ci, va, int(cuts[i]), int(cuts[i + 1]), int(cuts[j]), int(cuts[j + 1]))
            b = blocks[i][j]
            assert np.array_equal(b.rowptr, brp) and np.array_equal(b.colidx, bci)
            assert np.array_equal(b.vals.view(np.uint32), bva.view(np.uint32))
            assert b.m == int(cuts[j + 1] - cuts[j])
            assert np.array_equal(only[j].colidx, bci)
            total += b.nnz
    assert total == A.nnz


@pytest.mark.parametrize("P", [1, 2, 3, 8])
def test_needed_rows_match_compute_comm(mg, P):
    """mggcn_csr_needed_rows against prep.py's compute_comm loop restated in numpy
    (test/data/prep.py:262-268: mask the columns of block row i, count the mask inside every column block)."""
    rng = np.random.default_rng(10 + P)
    n = 517
    rp, ci, va = random_csr(rng, n, n, 0.03)
    A = mg.Csr(rp, ci, va, n)
    cuts = mg.row_cuts(n, P)
    L = A.needed_rows(cuts, cuts)
    want = np.zeros((P, P), dtype=np.uint64)
    for i in range(P):
        mask = np.zeros(n, dtype=bool)
        mask[ci[int(rp[cuts[i]]):int(rp[cuts[i + 1]])]] = True
        for j in range(P):
            want[i, j] = mask[int(cuts[j]):int(cuts[j + 1])].sum()
    assert np.array_equal(L, want)
    # never more than the block holds, and the blocks of a row together hold every distinct column of that row block
    assert (L <= np.diff(cuts)[None, :]).all()
    blocks = A.block_rows(cuts, cuts)
    for i in range(P):
        for j in range(P):
            assert int(L[i, j]) == len(np.unique(blocks[i][j].colidx))
    # an empty matrix needs nothing; a column outside the matrix is refused
    E = mg.Csr(np.zeros(n + 1, np.uint64), np.zeros(0, np.uint32), np.zeros(0, np.float32), n)
    assert not E.needed_rows(cuts, cuts).any()
    bad = mg.Csr(np.array([0, 1], np.uint64), np.array([9], np.uint32), np.ones(1, np.float32), 4)
    with pytest.raises(mg.MggcnError):
        bad.needed_rows(np.array([0, 1], np.uint32), np.array([0, 4], np.uint32))


def test_generator_emulates_prep(mg):
    """prep.py:101-126: N, F padded to x8, one self-loop per node, unit weights, symmetric pattern;
    padding rows are isolated (self-loop only), zero features, label 0."""
    shape = (1001, 1001 * 12, 13, 7, 1, 2.5, 200)
    A, X, Y, C = mg.generate(shape, seed=5, permute=True)
    assert A.n == 1008 and A.m == 1008 and X.shape == (1008, 16) and Y.shape == (1008,)
    assert np.all(A.vals == 1.0)
    rows = np.repeat(np.arange(A.n), np.diff(A.rowptr.astype(np.int64)))
    cols = A.colidx.astype(np.int64)
    assert np.all(np.diff(A.rowptr.astype(np.int64)) >= 1)
    key = rows * A.n + cols
    assert len(np.unique(key)) == len(key)                      # no duplicate entries
    assert np.array_equal(np.sort(key), np.sort(cols * A.n + rows))  # symmetric
    assert np.count_nonzero(rows == cols) == A.n                # exactly one self-loop per node
    for r in range(A.n):                                        # sorted rows
        seg = cols[int(A.rowptr[r]):int(A.rowptr[r + 1])]
        assert np.all(np.diff(seg) > 0)
    deg = np.diff(A.rowptr.astype(np.int64))
    pad = np.nonzero((np.abs(X).sum(1) == 0) & (deg == 1) & (Y == 0))[0]
    assert len(pad) >= 7
    assert Y.min() >= 0 and Y.max() < C and np.all(X[:, 13:] == 0)
    # deterministic, independent of thread count
    os.environ["OMP_NUM_THREADS"] = "1"
    A2, X2, Y2, _ = mg.generate(shape, seed=5, permute=True)
    assert np.array_equal(A.colidx, A2.colidx) and np.array_equal(X, X2) and np.array_equal(Y, Y2)
    A3, _, _, _ = mg.generate(shape, seed=6, permute=True)
    assert not np.array_equal(A.rowptr, A3.rowptr)


def test_dataset_directory_layout(mg, tmp_path):
    A, X, Y, C = mg.generate("oracle4k", seed=1, scale=0.05)
    mg.write_dataset(str(tmp_path / "ds"), A, X, Y)
    for f in ("graph.bin", "features.bin", "labels.bin", "sets.bin"):
        assert (tmp_path / "ds" / f).exists()
    B, info = mg.read_csr(str(tmp_path / "ds" / "graph.bin"))
    assert np.array_equal(B.colidx, A.colidx)
    assert np.array_equal(mg.read_dense(str(tmp_path / "ds" / "features.bin")), X)
    assert np.array_equal(mg.read_dense(str(tmp_path / "ds" / "labels.bin"), np.int32).ravel(), Y)


def test_init_weights_match_reference_stream(mg, orc):
    import json
    ka = json.load(open(os.path.join(GOLD, "known_answers.json")))
    assert np.array_equal(mg.init_weights(4, 3).ravel(), np.array(ka["init"]["W_4x3_default_gain"], np.float32))
    assert np.array_equal(mg.init_weights(608, 128), orc.init_weights(608, 128))


def test_class_padding_rule(mg):
    assert [mg.pad_classes(41, P) for P in (1, 2, 4, 8)] == [41, 42, 44, 48]  # src/main.cpp:135


@pytest.mark.parametrize("shape,scale,P", [("oracle4k", 1.0, 3), ("reddit", 0.01, 2), ("proteins", 0.02, 5),
                                           ("papers100M", 0.00004, 8)])
def test_sharded_generation_and_prep_bit_exact(mg, shape, scale, P):
    """N4: a rank builds ONLY its row slab (mggcn_gen_rows_begin) and derives both aggregation operands from the
    slab plus the vertex degrees; the P x P blocks must equal normalise -> transpose -> split of the whole graph
    (src/main.cpp:139-151) bit for bit, also when N % P != 0."""
    permute = shape != "papers100M"
    A, X, Y, C = mg.generate(shape, seed=1, scale=scale, permute=permute)
    N = A.n
    cuts = mg.row_cuts(N, P)
    deg = np.diff(A.rowptr.astype(np.int64)).astype(np.uint32)
    A2 = A.copy()
    A2.normalize(True)
    AT = A2.transpose()
    for r in range(P):
        r0, r1 = int(cuts[r]), int(cuts[r + 1])
        slab, Xl, Yl, _ = mg.generate_rows(shape, r0, r1, seed=1, scale=scale, permute=permute)
        assert np.array_equal(slab.rowptr, A.rowptr[r0:r1 + 1] - A.rowptr[r0])
        assert np.array_equal(slab.colidx, A.colidx[int(A.rowptr[r0]):int(A.rowptr[r1])])
        assert np.array_equal(Xl, X[r0:r1]) and np.array_equal(Yl, Y[r0:r1])
        fwd, bwd = mg.slab_operands(slab, deg, r0)
        local = np.array([0, slab.n], np.uint32)
        for mine, want in ((fwd.block_rows(local, cuts, only_row=0), AT.block_rows(cuts, cuts, only_row=r)),
                           (bwd.block_rows(local, cuts, only_row=0), A2.block_rows(cuts, cuts, only_row=r))):
            for a, b in zip(mine, want):
                assert np.array_equal(a.rowptr, b.rowptr) and np.array_equal(a.colidx, b.colidx)
                assert np.array_equal(a.vals.view(np.uint32), b.vals.view(np.uint32))
