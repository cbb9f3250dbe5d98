"""The allocation-free BatchNorm-statistics entry point, driven through the C ABI.

`tdl_bn_fwd_stats_noalloc` exists for callers that cannot give the library scratch memory because they cannot
get any themselves: processes whose device memory was handed in full to their own arena or to a serving
runtime (a late cudaMalloc fails or is against policy), and latency-critical loops where cudaMalloc/cudaFree
would stall the whole device.  It keeps its per-CTA partial sums in a fixed 4 MiB array inside the library.

This file is named so that pytest collects it last: a run that stops at the first failure has then already
executed every other test.  The test makes two calls back to back (second call on a different shape, so stale
partials from the first would show up) and compares each with the ordinary entry point that takes caller scratch.
Every buffer is created before the first native call; the calls themselves go through ctypes only.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

_i, _i64, _f = ctypes.c_int, ctypes.c_int64, ctypes.c_float


def _p(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _rel(a, b):
    return float((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6))


def _make(rows, C, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.randn(rows, C, device="cuda", generator=g) * 2 + 0.5).bfloat16()
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.1
    return x, gamma, beta


def _buffers(C):
    return {"rm": torch.zeros(C, device="cuda"), "rv": torch.ones(C, device="cuda"),
            "out": torch.zeros(4, C, device="cuda")}


def test_noalloc_matches_scratch_entry_point_twice_in_a_row():
    from tensorflowdistributedlearning_b200 import _native

    lib = _native.lib()
    lib.tdl_bn_scratch_floats.restype = ctypes.c_int64
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    cases = [(6000, 256, 11), (1234, 64, 12)]  # second call: different rows/C than the first
    prepared = []
    for rows, C, seed in cases:  # allocate everything up front
        x, gamma, beta = _make(rows, C, seed)
        need = int(lib.tdl_bn_scratch_floats(_i64(rows), _i(C)))
        prepared.append((rows, C, x, gamma, beta, torch.empty(need, device="cuda"), _buffers(C), _buffers(C)))
    torch.cuda.synchronize()
    for rows, C, x, gamma, beta, scratch, ref, new in prepared:
        rc = lib.tdl_bn_fwd_stats(_p(x), _i64(rows), _i(C), _p(scratch), _i64(scratch.numel()), _p(gamma), _p(beta),
                                  _p(ref["rm"]), _p(ref["rv"]), _f(0.99), _f(1e-3), _p(ref["out"][0]),
                                  _p(ref["out"][1]), _p(ref["out"][2]), _p(ref["out"][3]), stream)
        assert rc == 0, lib.tdl_last_error()
        rc = lib.tdl_bn_fwd_stats_noalloc(_p(x), _i64(rows), _i(C), _p(gamma), _p(beta), _p(new["rm"]), _p(new["rv"]),
                                          _f(0.99), _f(1e-3), _p(new["out"][0]), _p(new["out"][1]), _p(new["out"][2]),
                                          _p(new["out"][3]), stream)
        assert rc == 0, lib.tdl_last_error()
        torch.cuda.synchronize()
        for k in range(4):  # mean, invstd, scale, shift
            assert _rel(new["out"][k], ref["out"][k]) < 1e-4, (rows, C, k)
        assert _rel(new["rm"], ref["rm"]) < 1e-4 and _rel(new["rv"], ref["rv"]) < 1e-4
        # and both agree with a plain fp32 PyTorch computation of the same statistics
        xf = x.float()
        assert _rel(new["out"][0], xf.mean(0)) < 1e-4
        assert _rel(1 / new["out"][1] ** 2 - 1e-3, xf.var(0, unbiased=False)) < 1e-3


def test_noalloc_rejects_shapes_beyond_the_builtin_array():
    from tensorflowdistributedlearning_b200 import _native

    lib = _native.lib()
    C = 1 << 20  # 2 * C partial floats > the 1,048,576-float array
    x = torch.zeros(8, C, device="cuda", dtype=torch.bfloat16)
    v = torch.zeros(C, device="cuda")
    out = torch.zeros(4, C, device="cuda")
    rc = lib.tdl_bn_fwd_stats_noalloc(_p(x), _i64(8), _i(C), _p(v), _p(v), _p(None), _p(None), _f(0.99), _f(1e-3),
                                      _p(out[0]), _p(out[1]), _p(out[2]), _p(out[3]),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.tdl_last_error().decode()
    assert "built-in array" in msg and "1048576" in msg
