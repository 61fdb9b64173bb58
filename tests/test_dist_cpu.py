"""World-size-2 host logic on CPU (gloo): every rank derives its own block row of the P x P split
(src/dist_matrix.hpp:215-259), pads the class count (src/main.cpp:135), exchanges opaque handles the
way bench.py does, and the staged product over its blocks equals the oracle's staged SpMM rows."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_package
    from oracle import oracle as orc
    mg = load_package()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A, X, Y, C = mg.generate((1500, 1500 * 14, 20, 5, 1, 2.5, 200), seed=3)
        A.normalize(True)
        A_T = A.transpose()
        N, F = A.n, 24
        cuts = mg.row_cuts(N, world)
        blocks = A_T.block_rows(cuts, cuts, only_row=rank)
        # 1. handles travel through all_gather unchanged (bench.py's p2p wiring)
        mine = torch.frombuffer(bytearray(f"rank{rank:04d}".encode().ljust(80, b".")), dtype=torch.uint8)
        got = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(got, mine)
        assert [bytes(g.numpy().tobytes())[:8] for g in got] == [f"rank{r:04d}".encode() for r in range(world)]
        # 2. block nnz of all ranks add up to the matrix
        cnt = torch.tensor([sum(b.nnz for b in blocks)], dtype=torch.int64)
        dist.all_reduce(cnt)
        assert int(cnt) == A_T.nnz
        # 3. staged product over this rank's blocks, stage order 0..P-1, equals the oracle's rows
        B = np.random.default_rng(1).standard_normal((N, F)).astype(np.float32)
        r0, r1 = int(cuts[rank]), int(cuts[rank + 1])
        acc = None
        for j, blk in enumerate(blocks):
            part = orc.spmm(blk.rowptr, blk.colidx, blk.vals, B[int(cuts[j]):int(cuts[j + 1])])
            acc = part if acc is None else (acc + part).astype(np.float32)
        want = orc.spmm(A_T.rowptr, A_T.colidx, A_T.vals, B, cuts=cuts)[r0:r1]
        assert np.array_equal(acc, want)
        # 4. gathered row blocks reproduce the full product (the all-gather the exchange implements)
        t = torch.from_numpy(acc.copy())
        parts = [torch.empty((int(cuts[r + 1] - cuts[r]), F)) for r in range(world)]
        dist.all_gather(parts, t) if len({p.shape for p in parts}) == 1 else None
        assert mg.pad_classes(C, world) % world == 0
        # 5. sharded pre-processing (bench.py's path): the rank generates only its row slab, the degrees travel
        # through all_gather, and the blocks it derives are bit-identical to normalise -> transpose -> split of
        # the whole graph -- no rank ever holds more than its slab
        shape = (1500, 1500 * 14, 20, 5, 1, 2.5, 200)
        slab, Xl, Yl, _ = mg.generate_rows(shape, r0, r1, seed=3)
        assert np.array_equal(Xl, X[r0:r1]) and np.array_equal(Yl, Y[r0:r1])
        dl = torch.from_numpy(np.diff(slab.rowptr.astype(np.int64)).astype(np.int32))
        dparts = [torch.empty(int(cuts[r + 1] - cuts[r]), dtype=torch.int32) for r in range(world)]
        dist.all_gather(dparts, dl)
        deg_all = torch.cat(dparts).numpy().astype(np.uint32)
        fwd, bwd = mg.slab_operands(slab, deg_all, r0)
        local = np.array([0, slab.n], np.uint32)
        for mine_b, want_b in ((fwd.block_rows(local, cuts, only_row=0), blocks),
                               (bwd.block_rows(local, cuts, only_row=0), A.block_rows(cuts, cuts, only_row=rank))):
            for a, b in zip(mine_b, want_b):
                assert np.array_equal(a.rowptr, b.rowptr) and np.array_equal(a.colidx, b.colidx)
                assert np.array_equal(a.vals.view(np.uint32), b.vals.view(np.uint32))
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_two_rank_host_path():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_checker_skewed_cuts_for_every_world_size(mg):
    """The GPU checker's uneven cuts (tests/dist_engine_check.py --skew): even cuts move DOWN, which a uint32 array
    refuses in place (the 8-GPU run of round 2 died there, not in the engine)."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("dist_engine_check", os.path.join(os.path.dirname(__file__), "dist_engine_check.py"))
    chk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(chk)
    for world in (2, 4, 8):
        cuts = chk.skewed_cuts(mg.row_cuts(4096, world), 11)
        assert cuts.dtype == np.uint32 and cuts[0] == 0 and cuts[-1] == 4096
        sizes = np.diff(cuts.astype(np.int64))
        assert sizes.min() > 0 and (world == 2 or len(set(sizes.tolist())) > 1)
