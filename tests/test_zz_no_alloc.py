"""pb200_dot_noalloc -- the entry point for callers that cannot allocate device memory (their GPU memory
belongs to their own pool or framework allocator, or they sit in a loop that cannot tolerate the device-wide
synchronisation of cudaMalloc/cudaFree). Its partial sums live in a fixed array inside the library image
instead of a cudaMalloc'ed buffer, which is why its size is capped and why all calls on a GPU share one scratch.

Checked here, through the C ABI, against the existing allocating entry point pb200_op_dot on the same data:
two calls in a row (the second reuses the scratch the first one dirtied), the size cap, and a call on a
non-default stream. The file sorts last so that a run stopping at the first failure has already run everything
else; it must also pass alone in a fresh process (the first call on a device does the address look-up).
"""
import numpy as np
import pytest

from __graft_entry__ import load_package
from oracle_lib import seeded_field

pb = load_package()
pytestmark = pytest.mark.gpu

RTOL_DOT = 1e-13   # the tolerance tests/test_gpu_parity.py uses for inner products


def to_device_block(torch, host, pitch):
    """(nx+2, ny+2) host array -> device tensor in the solver's layout (rows of `pitch` doubles)."""
    dev = torch.zeros((host.shape[0], pitch), dtype=torch.float64, device="cuda")
    dev[:, :host.shape[1]] = torch.from_numpy(host).cuda()
    return dev


def noalloc_dot(torch, u, v, nx, ny, h1, h2, stream=None):
    pitch = ((ny + 2 + 15) // 16) * 16
    du, dv = to_device_block(torch, u, pitch), to_device_block(torch, v, pitch)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    pb.dot_noalloc(du.data_ptr(), dv.data_ptr(), nx, ny, pitch, h1, h2, out.data_ptr(),
                   None if stream is None else stream.cuda_stream)
    torch.cuda.synchronize()
    return float(out.item())


def test_two_calls_in_a_row_match_the_allocating_entry_point(oracle):
    import torch
    for nx, ny, seed in [(300, 513, 1), (300, 513, 2), (39, 39, 3), (1000, 64, 4)]:
        h1, h2 = 2.0 / (nx + 1), 1.2 / (ny + 1)
        u, v = seeded_field(nx, ny, seed=seed), seeded_field(nx, ny, seed=seed + 100)
        want = pb.op_dot(pb.default_params(nx + 1, ny + 1), u, v)
        got = noalloc_dot(torch, u, v, nx, ny, h1, h2)
        assert abs(got - want) <= RTOL_DOT * abs(want), (nx, ny, seed, got, want)
        cpu = oracle.dot(u, v, h1, h2)          # and against the CPU oracle directly, not only via pb200_op_dot
        assert abs(got - cpu) <= RTOL_DOT * abs(cpu), (nx, ny, seed, got, cpu)


def test_same_scratch_on_another_stream_after_the_first_call():
    import torch
    nx, ny = 257, 300
    h1, h2 = 2.0 / (nx + 1), 1.2 / (ny + 1)
    u, v = seeded_field(nx, ny, seed=9), seeded_field(nx, ny, seed=10)
    want = pb.op_dot(pb.default_params(nx + 1, ny + 1), u, v)
    s = torch.cuda.Stream()
    for _ in range(2):
        got = noalloc_dot(torch, u, v, nx, ny, h1, h2, stream=s)
        assert abs(got - want) <= RTOL_DOT * abs(want)


def test_size_cap_is_reported_not_overrun():
    import torch
    t = torch.zeros(16, dtype=torch.float64, device="cuda")
    ny = 256 * 17                       # 17 column blocks x 512 row blocks = 8704 > 8192
    with pytest.raises(pb.PB200Error, match="reduction blocks"):
        pb.dot_noalloc(t.data_ptr(), t.data_ptr(), 512, ny, ny + 2, 1.0, 1.0, t.data_ptr())
    with pytest.raises(pb.PB200Error, match="pitch"):
        pb.dot_noalloc(t.data_ptr(), t.data_ptr(), 4, 4, 4, 1.0, 1.0, t.data_ptr())
