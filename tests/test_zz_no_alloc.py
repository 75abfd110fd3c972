"""dls_sq_quant_noalloc, called the way a non-torch host would: straight through the C ABI.

This variant of the stochastic quantizer exists for callers that must not allocate device
memory -- a C-ABI host whose HBM is already fully committed (a late cudaMalloc could fail),
or a worker thread sharing the GPU with many other streams (cudaMalloc/cudaFree stall all
of them).  Its scratch is an array inside the library, shared by every call, so the test
makes two calls in a row with different data: the second must not see the first one's
maxima.  Each result is compared with the ordinary entry point (caller-provided scratch)
run with the same seed; both run the same kernels, so the outputs must be identical.

The file name sorts last on purpose: a run that stops at the first failure has then
already exercised every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

NUMELS = [1728, 64, 64, 36864, 10, 7, 1, 5120, 33, 4099]


def _lib():
    from distributed_learning_simulator_b200.ops import runtime

    lib = ctypes.CDLL(runtime.lib_path())
    lib.dls_last_error.restype = ctypes.c_char_p
    vp, ll, i32, u64 = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_uint64
    lib.dls_sq_quant.argtypes = [vp, ll, vp, i32, i32, u64, vp, vp, vp, vp, vp]
    lib.dls_sq_quant.restype = i32
    lib.dls_sq_quant_noalloc.argtypes = [vp, ll, vp, i32, i32, u64, vp, vp, vp, vp]
    lib.dls_sq_quant_noalloc.restype = i32
    lib.dls_sq_quant_noalloc_max_tensors.restype = i32
    return lib


def _quantize(lib, flat, offsets, n_tensors, seed, no_alloc):
    dev = flat.device
    n = flat.numel()
    payload = torch.zeros(n, dtype=torch.uint8, device=dev)
    sign = torch.zeros((n + 7) // 8, dtype=torch.uint8, device=dev)
    scale = torch.zeros(n_tensors, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    if no_alloc:
        rc = lib.dls_sq_quant_noalloc(flat.data_ptr(), n, offsets.data_ptr(), n_tensors, 255, seed,
                                      payload.data_ptr(), sign.data_ptr(), scale.data_ptr(), stream)
    else:
        scratch = torch.empty(n_tensors, dtype=torch.int32, device=dev)
        rc = lib.dls_sq_quant(flat.data_ptr(), n, offsets.data_ptr(), n_tensors, 255, seed, payload.data_ptr(),
                              sign.data_ptr(), scale.data_ptr(), scratch.data_ptr(), stream)
    assert rc == 0, lib.dls_last_error().decode()
    torch.cuda.synchronize(dev)
    return payload, sign, scale


def test_no_alloc_entry_point_matches_regular_one_twice():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    offs = [0]
    for k in NUMELS:
        offs.append(offs[-1] + k)
    offsets = torch.tensor(offs, dtype=torch.int64, device=dev)
    g = torch.Generator(device="cuda").manual_seed(5)
    # second call uses much smaller values: stale maxima in the shared scratch would show
    for call, magnitude in enumerate((10.0, 1e-3)):
        flat = torch.randn(offs[-1], device=dev, generator=g) * magnitude
        ref = _quantize(lib, flat, offsets, len(NUMELS), seed=77 + call, no_alloc=False)
        got = _quantize(lib, flat, offsets, len(NUMELS), seed=77 + call, no_alloc=True)
        torch.testing.assert_close(got[2], ref[2], rtol=0, atol=0)
        assert torch.equal(got[0], ref[0]), f"payload differs on call {call}"
        assert torch.equal(got[1], ref[1]), f"sign plane differs on call {call}"
        for t in range(len(NUMELS)):
            expect = flat[offs[t]:offs[t + 1]].abs().max() / 255
            torch.testing.assert_close(got[2][t], expect, rtol=1e-6, atol=0)


def test_no_alloc_entry_point_rejects_oversized_models():
    lib = _lib()
    dev = torch.device("cuda", 0)
    limit = lib.dls_sq_quant_noalloc_max_tensors()
    assert limit == 4096
    n_tensors = limit + 1
    flat = torch.ones(n_tensors, device=dev)
    offsets = torch.arange(n_tensors + 1, dtype=torch.int64, device=dev)
    out = torch.zeros(n_tensors, dtype=torch.uint8, device=dev)
    scale = torch.zeros(n_tensors, dtype=torch.float32, device=dev)
    rc = lib.dls_sq_quant_noalloc(flat.data_ptr(), n_tensors, offsets.data_ptr(), n_tensors, 255, 1, out.data_ptr(),
                                  out.data_ptr(), scale.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    assert rc != 0
    assert "exceed" in lib.dls_last_error().decode()
    torch.cuda.synchronize(dev)
    assert int(out.sum().item()) == 0  # nothing ran
