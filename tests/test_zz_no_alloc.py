"""GPU: as_inclusive_scan_u32_static -- the prefix sum for callers that cannot allocate.

Some hosts of this library have no allocator to lend it: a robot's mapping process that divided all
of its device memory into fixed arenas at start-up, or a loop with a latency budget where
cudaMalloc / cudaFree are forbidden because they may stall the device.  For them the scan's
look-back state sits in a fixed array inside the library (at most 4,194,304 elements per call).

The test calls the entry point through the C ABI with raw device pointers and checks it against the
ordinary entry point (caller-provided scratch): integers, so the results must be identical.  It does
so on two calls in a row of different sizes -- the second one finds the first call's leftovers in the
shared array and must still be right -- and checks that an oversized request is refused.
This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.
"""
import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (sys.path)

pytestmark = pytest.mark.gpu


def _scan_both(L, capi, x, stream):
    n = x.size
    dev = "cuda:0"
    x_d = torch.from_numpy(x.view(np.int32)).to(dev)
    ref_out = torch.empty_like(x_d)
    ref_tot = torch.zeros(1, dtype=torch.int32, device=dev)
    tb = L.as_scan_temp_bytes(n)
    temp = torch.empty(tb, dtype=torch.uint8, device=dev)
    capi.check(L.as_inclusive_scan_u32(x_d.data_ptr(), ref_out.data_ptr(), n, ref_tot.data_ptr(), temp.data_ptr(), tb,
                                       stream))
    out = torch.empty_like(x_d)
    tot = torch.zeros(1, dtype=torch.int32, device=dev)
    capi.check(L.as_inclusive_scan_u32_static(x_d.data_ptr(), out.data_ptr(), n, tot.data_ptr(), stream))
    torch.cuda.synchronize()
    return (out.cpu().numpy().view(np.uint32), tot.cpu().numpy().view(np.uint32),
            ref_out.cpu().numpy().view(np.uint32), ref_tot.cpu().numpy().view(np.uint32))


def test_static_scratch_scan_matches_regular_entry_point_twice_in_a_row():
    from activesplat_b200 import capi
    L = capi.lib()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    assert L.as_static_scan_max_n() == 4_194_304
    rng = np.random.default_rng(2024)
    # large first (fills many tile states), then small (must not be confused by what is left behind),
    # then the largest size the fixed array allows
    for n in (1_000_003, 5_000, 4_194_304):
        x = rng.integers(0, 40, n, dtype=np.uint32)
        out, tot, ref_out, ref_tot = _scan_both(L, capi, x, stream)
        np.testing.assert_array_equal(out, ref_out)
        np.testing.assert_array_equal(tot, ref_tot)
        np.testing.assert_array_equal(out, np.cumsum(x.astype(np.uint64)).astype(np.uint32))


def test_static_scratch_scan_refuses_oversized_input():
    from activesplat_b200 import capi
    L = capi.lib()
    torch.cuda.set_device(0)
    n = L.as_static_scan_max_n() + 1
    x = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    y = torch.full((n,), 7, dtype=torch.int32, device="cuda:0")
    rc = L.as_inclusive_scan_u32_static(x.data_ptr(), y.data_ptr(), n, None, torch.cuda.current_stream().cuda_stream)
    assert rc == capi.AS_ERR_STATIC_LIMIT
    msg = L.as_error_string(rc).decode()
    assert "4194304" in msg and "as_inclusive_scan_u32" in msg
    torch.cuda.synchronize()
    assert int(y[0].item()) == 7 and int(y[-1].item()) == 7  # nothing was launched
    with pytest.raises(capi.AsError, match="at most 4194304 elements"):
        capi.check(rc)
