"""gs_bin_scan_static: the tile-count scan for callers that cannot allocate device memory.

Every other scan entry point of the C ABI takes its look-back scratch from the caller. That suits torch, but
not a host that is only ever handed input and output buffers (a custom-op slot of an inference server, a cgo /
JNI binding whose runtime owns the device memory) or one that sits where cudaMalloc is forbidden because it
synchronises the device. For those the library carries its own scratch, a 64 KiB `__device__` array, behind
`gs_bin_scan_static`; the price is a size limit (GS_SCAN_STATIC_MAX_N elements) and one shared array per GPU.

What is checked here, through the C ABI only:
  * the results equal gs_bin_scan's bit for bit (integer work: the tolerance of the binning tests is zero),
    on two calls in a row — the second call finds the first one's ticket and status words in the array and must
    not be disturbed by them, whether it is shorter, longer or the same length;
  * a call does not lower the free device memory (it really allocates nothing);
  * sizes above the limit are refused with GS_ERR_UNSUPPORTED before any CUDA call is made.

This file is named test_zz_* so that pytest collects it after every other file: under `-x` a failure here cannot hide
another test. It does not depend on anything the earlier tests leave behind (and passes alone in a fresh process):
the first call on the device resolves the array's address, any later call reuses it.
"""
import re
import os

import pytest
import torch

from gaussctrl_exp_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_constant(name: str) -> int:
    hdr = open(os.path.join(ROOT, "include", "gs_b200.h")).read()
    expr = re.search(rf"#define\s+{name}\s+(.+)", hdr).group(1)
    expr = re.sub(r"\b(GS_[A-Z_]+)\b", lambda m: str(_header_constant(m.group(1))), expr).replace("/", "//")
    return int(eval(expr, {"__builtins__": {}}))


MAX_N = _header_constant("GS_SCAN_STATIC_MAX_N")


def _scan_with_caller_scratch(L, hit):
    """The existing entry point: scratch comes from the caller (here: from torch)."""
    n = hit.numel()
    cum = torch.empty_like(hit)
    total = torch.full((1,), -7, device=hit.device, dtype=torch.int32)
    scratch = torch.empty(L.gs_scan_scratch_bytes(n) + 256, device=hit.device, dtype=torch.uint8)
    _lib.check(L.gs_bin_scan(n, _lib.ptr(hit), _lib.ptr(cum), _lib.ptr(total), _lib.ptr(scratch), scratch.numel(),
                             _lib.stream_ptr()), "gs_bin_scan")
    return cum, total


def _scan_without_scratch(L, hit, cum, total):
    """The entry point under test: inputs and outputs only, preallocated by the caller."""
    _lib.check(L.gs_bin_scan_static(hit.numel(), _lib.ptr(hit), _lib.ptr(cum), _lib.ptr(total), _lib.stream_ptr()),
               "gs_bin_scan_static")


def test_limit_is_enforced_without_a_gpu():
    """N above the documented maximum is an error, not a truncated scan; the check comes before any CUDA call."""
    L = _lib.lib()
    assert MAX_N == 16 * 1024 * 1024
    assert _header_constant("GS_SCAN_STATIC_BYTES") == 16 + 8 * (MAX_N // 2048 + 1) <= 65 * 1024
    assert L.gs_scan_scratch_bytes(MAX_N) <= _header_constant("GS_SCAN_STATIC_BYTES")
    assert L.gs_bin_scan_static(MAX_N + 1, None, None, None, None) == -3      # GS_ERR_UNSUPPORTED
    assert L.gs_bin_scan_static(-1, None, None, None, None) == -1             # GS_ERR_BAD_ARG
    with pytest.raises(_lib.GsError, match="GS_ERR_UNSUPPORTED"):
        _lib.check(L.gs_bin_scan_static(2**31 - 1, None, None, None, None), "gs_bin_scan_static")


@pytest.mark.gpu
def test_two_calls_in_a_row_match_gs_bin_scan(cuda_dev):
    L = _lib.lib()
    g = torch.Generator(device="cpu").manual_seed(20261106)
    # long, then short (stale status words beyond the short call's tiles), then long again, tile-edge sizes, the limit
    sizes = [1_000_003, 1_000_003, 5, 2048, 2049, 300_000, 4_100_000, 1, MAX_N]
    with torch.cuda.device(cuda_dev):
        for n in sizes:
            hit = torch.randint(0, 9, (n,), generator=g, dtype=torch.int32).to(cuda_dev)
            want_cum, want_total = _scan_with_caller_scratch(L, hit)
            got_cum = torch.empty_like(hit)
            got_total = torch.full((1,), -9, device=cuda_dev, dtype=torch.int32)
            for call in range(2):                                   # two calls in a row on the shared array
                got_cum.fill_(-1)
                got_total.fill_(-9)
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info()[0]
                _scan_without_scratch(L, hit, got_cum, got_total)
                torch.cuda.synchronize()
                assert torch.cuda.mem_get_info()[0] >= free_before or call == 0, "the call took device memory"
                assert torch.equal(got_cum, want_cum), f"N={n} call {call}: cumulative counts differ"
                assert int(got_total.item()) == int(want_total.item()) == int(hit.sum(dtype=torch.int64).item())
        # back-to-back on one stream with no host synchronisation in between, different inputs
        a = torch.randint(0, 50, (700_001,), generator=g, dtype=torch.int32).to(cuda_dev)
        b = torch.randint(0, 3, (123_457,), generator=g, dtype=torch.int32).to(cuda_dev)
        ca, ta = torch.empty_like(a), torch.zeros(1, device=cuda_dev, dtype=torch.int32)
        cb, tb = torch.empty_like(b), torch.zeros(1, device=cuda_dev, dtype=torch.int32)
        _scan_without_scratch(L, a, ca, ta)
        _scan_without_scratch(L, b, cb, tb)
        torch.cuda.synchronize()
        for hit, cum, tot in ((a, ca, ta), (b, cb, tb)):
            want_cum, want_total = _scan_with_caller_scratch(L, hit)
            assert torch.equal(cum, want_cum) and torch.equal(tot, want_total)
        # N = 0: only the total is written
        t0 = torch.full((1,), 5, device=cuda_dev, dtype=torch.int32)
        _lib.check(L.gs_bin_scan_static(0, None, None, _lib.ptr(t0), _lib.stream_ptr()), "gs_bin_scan_static")
        assert int(t0.item()) == 0
