"""mggcn_colsum_static_f32: column sums for callers that have no way to get device memory.

Some users of libmggcn.so cannot hand it a workspace: a plug-in inside an application whose own memory
manager holds the whole GPU, code running where cudaMalloc is not allowed (a capture region, a host
callback), or a loop that cannot afford the device-wide stall an allocation brings.  For them the library
carries a fixed scratch array (592 x 1024 floats per GPU) and this entry point, which takes only the input
and the output and allocates nothing.

The file name starts with `zz` on purpose: pytest collects it last, so with `-x` a failure here cannot hide
any other test.  It also has to pass when it is the only file run in a fresh process -- the first call on a
device is the one that resolves the scratch array's address -- which is why nothing here depends on state
left behind by other tests.
"""
import os
import re

import numpy as np
import pytest

# tests/test_gpu_dense.py::test_colsum_and_elementwise holds mggcn_colsum_f32 to this absolute error
COLSUM_ATOL = 5e-4
MAX_N = 1024


def _static_colsum(mg, torch, G):
    """One call through the C-ABI.  torch owns the input and the output tensor, nothing else."""
    rows, N = G.shape
    out = torch.full((N,), float("nan"), dtype=torch.float32, device=G.device)
    rc = mg.lib().mggcn_colsum_static_f32(mg._stream(), rows, N, G.data_ptr(), G.stride(0), out.data_ptr())
    assert rc == 0, mg.lib().mggcn_last_error()
    return out


@pytest.mark.gpu
def test_static_scratch_colsum_matches_workspace_colsum_twice(mg, cuda):
    """Two calls back to back, the second on a smaller and narrower matrix.

    The second call finds the first call's partial sums still in the shared array (more partial rows, a
    different row pitch), so a kernel that read anything it had not just written would show up here.  Both
    results are compared with the existing workspace entry point mggcn_colsum_f32 on the same input.
    """
    torch = cuda
    rng = np.random.default_rng(31)
    shapes = [(40_003, 257), (1_501, 48)]
    for rows, N in shapes:
        G = torch.from_numpy((rng.standard_normal((rows, N)) * 3 + 50).astype(np.float32)).cuda()
        want = mg.colsum(G)                      # caller-allocated workspace
        got = _static_colsum(mg, torch, G)       # library-owned scratch
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=0, atol=COLSUM_ATOL)


@pytest.mark.gpu
def test_static_scratch_colsum_edges(mg, cuda):
    """The widest accepted input, an input wider than the limit, an unaligned view and an empty one."""
    torch = cuda
    rng = np.random.default_rng(32)
    G = torch.from_numpy(rng.standard_normal((700, MAX_N)).astype(np.float32)).cuda()
    np.testing.assert_allclose(_static_colsum(mg, torch, G).cpu().numpy(), mg.colsum(G).cpu().numpy(),
                               rtol=0, atol=COLSUM_ATOL)
    # one column too many: refused before anything is launched, with a message that names the limit
    wide = torch.zeros((4, MAX_N + 1), dtype=torch.float32, device="cuda")
    out = torch.zeros(MAX_N + 1, dtype=torch.float32, device="cuda")
    before = mg.lib().mggcn_launch_count()
    rc = mg.lib().mggcn_colsum_static_f32(mg._stream(), 4, MAX_N + 1, wide.data_ptr(), wide.stride(0), out.data_ptr())
    assert rc == -1 and b"1024" in mg.lib().mggcn_last_error()
    assert mg.lib().mggcn_launch_count() == before
    # a view that starts one float into each row takes the scalar-load kernel instance
    base = torch.from_numpy(rng.standard_normal((333, 20)).astype(np.float32)).cuda()
    view = base[:, 1:18]
    np.testing.assert_allclose(_static_colsum(mg, torch, view).cpu().numpy(), mg.colsum(view).cpu().numpy(),
                               rtol=0, atol=COLSUM_ATOL)
    # no rows: every column sums to zero
    empty = torch.zeros((0, 16), dtype=torch.float32, device="cuda")
    assert torch.count_nonzero(_static_colsum(mg, torch, empty)).item() == 0


def test_static_scratch_entry_point_source_never_allocates(mg):
    """CPU-side guard: neither the entry point nor its address look-up names an allocating call."""
    src = open(os.path.join(os.path.dirname(mg.LIB_PATH), "..", "csrc", "dense.cu")).read()
    bodies = []
    for head in (r"int colsum_static_scratch\(float \*\*out\) \{", r"int mggcn_colsum_static_f32\([^)]*\) \{"):
        m = re.search(head, src)
        assert m, head
        end = src.index("\n}\n", m.end())
        bodies.append(src[m.end():end])
    text = "\n".join(bodies)
    assert "cudaGetSymbolAddress" in text and "g_colsum_scratch" in text
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMemPool", "new ", "malloc(", "torch", "at::"):
        assert banned not in text, banned
