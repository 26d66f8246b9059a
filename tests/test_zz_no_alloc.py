"""GPU test of b200_colsum_static, the column sum that allocates nothing.

Who needs it: hosts that cannot give the library scratch memory and must not have it allocate behind their
back — a minimal C binding without an allocator, a process whose framework pool already holds the GPU's memory
(a raw cudaMalloc there may fail or fragment the pool), or a step that overlaps communication and cannot afford
the device-wide synchronisation that cudaMalloc/cudaFree imply.  The entry point keeps its partial sums in a
1 MiB array inside the library instead.

This file sorts last on purpose: if it fails under `pytest -x`, every other test has already run.  It calls the
entry point through the C ABI twice in a row (the array keeps the previous call's contents, so the second call
must not be disturbed by them) and compares with b200_colsum, which uses caller-provided scratch.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_static(lib, dy, out, M, N, dtype_code):
    lib.colsum_static(dy.data_ptr(), dy.stride(0), out.data_ptr(), M, N, dtype_code,
                      torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32], ids=["bf16", "f32"])
def test_colsum_static_matches_colsum_twice_in_a_row(dtype):
    from b200step import ops
    from b200step.lib import get_lib

    lib = get_lib()
    code = ops.dtype_code(dtype)
    g = torch.Generator().manual_seed(7)
    cases = [(65536, 512), (4096, 1000)]   # two calls in a row, different shapes: stale partials must not leak
    for M, N in cases:
        dy = torch.randn(M, N, generator=g).to(dtype).cuda()
        ref = torch.full((N,), float("nan"), device="cuda")
        scratch = torch.empty(ops.colsum_scratch_floats(N), device="cuda")
        ops.colsum(dy, ref, scratch, M=M, N=N)
        got = torch.full((N,), float("nan"), device="cuda")
        _call_static(lib, dy, got, M, N, code)
        torch.cuda.synchronize()
        # same tolerance as test_colsum in test_kernels_gpu.py (fp32 partial sums, grouping may differ)
        torch.testing.assert_close(got, ref, rtol=1e-4, atol=1e-3 * (M ** 0.5) * 1e-2)
        # and again on the same input: identical to the first static call, bit for bit
        again = torch.full((N,), float("nan"), device="cuda")
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()      # the driver's own count of free device memory
        reserved_before = torch.cuda.memory_reserved()  # and what torch's pool holds
        _call_static(lib, dy, again, M, N, code)
        torch.cuda.synchronize()
        # the property the entry point exists for: the call took no device memory, from the driver or from torch
        assert torch.cuda.mem_get_info()[0] == free_before
        assert torch.cuda.memory_reserved() == reserved_before
        assert torch.equal(got, again)


def test_colsum_static_rejects_misaligned_and_null_pointers():
    """128-bit loads need a 16-byte aligned matrix: a misaligned view must be refused up front with an error
    message, not fault asynchronously on the device."""
    from b200step.lib import B200Error, get_lib

    lib = get_lib()
    base = torch.zeros(64 * 512 + 8, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(512, device="cuda")
    skew = base[1:1 + 64 * 512].view(64, 512)          # starts 2 bytes off a 16-byte boundary
    with pytest.raises(B200Error, match="16-byte aligned"):
        _call_static(lib, skew, out, 64, 512, 1)
    with pytest.raises(B200Error, match="NULL"):
        lib.colsum_static(None, 512, out.data_ptr(), 64, 512, 1, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()


def test_colsum_static_rejects_rows_wider_than_its_limit():
    from b200step.lib import B200Error, get_lib

    lib = get_lib()
    dy = torch.zeros(8, 4104, dtype=torch.float32, device="cuda")
    out = torch.zeros(4104, device="cuda")
    with pytest.raises(B200Error, match="static scratch limit"):
        _call_static(lib, dy, out, 8, 4104, 0)
