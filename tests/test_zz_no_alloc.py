"""GPU: the allocation-free statistics entry point against the regular one.

Collected last on purpose (file name): should it fail, a run that stops at the first
failure has already exercised everything else.

`b200dp_bn_fwd_stats_noalloc` serves hosts that cannot give the library an allocator —
a process whose own pool owns all device memory and passes only inputs and outputs.  Its
scratch is one fixed array shared by all calls, so the checks here are: same numbers as
the entry point with caller-provided scratch, also on a second call with different data
(nothing stale survives), and a clear error beyond the stated channel limit.
"""
import pytest
import torch

from dbx_distributed_pytorch_examples_b200 import _lib

pytestmark = pytest.mark.gpu


def both(x, C):
    lib = _lib.lib()
    M = x.numel() // C
    code = _lib.dtype_code(x.dtype)
    s = _lib.current_stream()
    scratch = torch.zeros(int(lib.b200dp_bn_scratch_bytes(C)), dtype=torch.uint8, device="cuda")
    want = torch.empty(2 * C + 1, device="cuda")
    got = torch.full((2 * C + 1,), float("nan"), device="cuda")
    _lib.check(lib.b200dp_bn_fwd_stats(x.data_ptr(), code, M, C, scratch.data_ptr(), want.data_ptr(), None, 0, s))
    _lib.check(lib.b200dp_bn_fwd_stats_noalloc(x.data_ptr(), code, M, C, got.data_ptr(), s))
    torch.cuda.synchronize()
    return got, want


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,C", [(64 * 56 * 56, 64), (4 * 7 * 7, 2048), (1000, 72), (3, 8)])
def test_matches_regular_entry_point_twice_in_a_row(dtype, M, C):
    torch.manual_seed(M + C)
    for call in range(2):  # the second call meets the first call's leftovers in the shared array
        x = (torch.randn(M, C, device="cuda") * (1.0 + call) + 0.3 * call).to(dtype)
        got, want = both(x, C)
        assert float(got[2 * C]) == float(want[2 * C]) == M
        # same kernel, same grid, same merge order: the tolerance of the other BN tests is ample
        torch.testing.assert_close(got[:C], want[:C], rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(got[C:2 * C], want[C:2 * C], rtol=1e-4, atol=1e-4)
        ref_mean = x.float().mean(0)
        torch.testing.assert_close(got[:C], ref_mean, rtol=1e-4, atol=1e-4)


def test_rejects_sizes_beyond_the_fixed_scratch():
    lib = _lib.lib()
    x = torch.zeros(8, 8, device="cuda")
    out = torch.empty(17, device="cuda")
    rc = lib.b200dp_bn_fwd_stats_noalloc(x.data_ptr(), _lib.F32, 2, 4096 + 8, out.data_ptr(), _lib.current_stream())
    assert rc == -5 and b"limit" in lib.b200dp_error_string(rc)
    assert lib.b200dp_bn_fwd_stats_noalloc(x.data_ptr(), _lib.F32, 8, 8, None, _lib.current_stream()) == -1
