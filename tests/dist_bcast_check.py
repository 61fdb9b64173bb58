"""Run under torchrun: mggcn_bcast_stage_f32 (row a5, replaces dist_row_dn_matrix::bcast,
src/dist_matrix.hpp:458-467) on a communicator of its own -- root i's block lands in every rank's staging
buffer, for every root in turn, bit for bit.  The communicator is created here with ncclCommInitRank through
ctypes (the C-ABI takes a plain ncclComm_t, exactly what the reference's dist_context holds)."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


class UniqueId(C.Structure):
    _fields_ = [("internal", C.c_char * 128)]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mg = load_package()
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = (C.c_char * 128)()
        mg.check(mg.lib().mggcn_nccl_unique_id(buf))
        idt.copy_(torch.frombuffer(bytearray(bytes(buf)), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    uid = UniqueId()
    C.memmove(C.byref(uid), bytes(idt.cpu().numpy().tobytes()), 128)
    nccl = C.CDLL("libnccl.so.2")  # the copy torch has already mapped
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, UniqueId, C.c_int]
    nccl.ncclCommInitRank.restype = C.c_int
    comm = C.c_void_p()
    rc = nccl.ncclCommInitRank(C.byref(comm), world, uid, rank)
    assert rc == 0, f"ncclCommInitRank -> {rc}"
    n, F = 1000 + 7 * rank, 44  # uneven blocks, class-width rows
    rng = np.random.default_rng(100 + rank)
    mine = torch.from_numpy(rng.standard_normal((n, F)).astype(np.float32)).cuda()
    st = torch.cuda.Stream()
    for root in range(world):
        rows = 1000 + 7 * root
        want = torch.from_numpy(np.random.default_rng(100 + root).standard_normal((rows, F)).astype(np.float32)).cuda()
        recv = torch.zeros((rows, F), device="cuda")
        torch.cuda.synchronize()
        mg.check(mg.lib().mggcn_bcast_stage_f32(comm, C.c_void_p(st.cuda_stream), mine.data_ptr(), recv.data_ptr(), rows * F, root))
        st.synchronize()
        assert torch.equal(recv, want), (rank, root)
    # error path: a null communicator is an argument error, not a crash
    assert mg.lib().mggcn_bcast_stage_f32(None, C.c_void_p(st.cuda_stream), mine.data_ptr(), mine.data_ptr(), 4, 0) == -1
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)
    dist.barrier()
    if rank == 0:
        print(f"bcast check ok: world={world}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
