"""grt_integrate_flux_noalloc (include/grt_sm100.h): the allocation-free variant of the
broadband integration, through the C ABI, against the existing grt_integrate_flux.

The entry point has no counterpart in the reference (added
for callers that cannot call cudaMalloc); its scratch is a file-scope
`static __device__` array that every call shares, so the test makes two calls in a row --
the second on fewer levels and other data, so that stale scratch contents would show -- and
checks the size limit.  Collected last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-12       # the driver / solver tests' tolerance on broadband sums (summation order)
MAX_LEVELS = 201   # GRT_NOALLOC_MAX_LEVELS = GRT_MAX_LAYERS + 1


def _abi(api):
    cu = api.cuda_layer()
    real_p = C.POINTER(api.REAL)
    cu.grt_malloc.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_size_t]
    cu.grt_free.argtypes = [C.c_int, C.c_void_p]
    cu.grt_copy_to_device.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
    cu.grt_copy_to_host.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]
    cu.grt_integrate_flux.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_double, C.c_void_p, real_p]
    cu.grt_integrate_flux_noalloc.argtypes = [C.c_int, C.c_int, C.c_uint64, C.c_double,
                                              C.c_void_p, C.c_void_p]
    cu.grt_last_error.argtypes = [C.c_char_p, C.c_int]
    return cu


def _both(api, cu, flux, dw, out_dev):
    """(existing entry point's sums, new entry point's sums) for flux[levels, n]."""
    levels, n = flux.shape
    size = flux.dtype.itemsize
    dev = C.c_void_p()
    assert cu.grt_malloc(0, C.byref(dev), flux.size*size) == 0
    try:
        assert cu.grt_copy_to_device(0, dev, flux.ctypes.data_as(C.c_void_p), flux.size*size) == 0
        want = np.empty(levels, dtype=api.DTYPE)
        assert cu.grt_integrate_flux(0, levels, n, dw, dev,
                                     want.ctypes.data_as(C.POINTER(api.REAL))) == 0
        assert cu.grt_integrate_flux_noalloc(0, levels, n, dw, dev, out_dev) == 0
        got = np.empty(levels, dtype=api.DTYPE)
        assert cu.grt_copy_to_host(0, got.ctypes.data_as(C.c_void_p), out_dev, levels*size) == 0
    finally:
        assert cu.grt_free(0, dev) == 0
    return want, got


def test_no_alloc_integration_matches_the_existing_entry_point_twice_in_a_row(built):
    api = built
    api.load()
    if api.num_gpus() < 1:
        pytest.skip("no CUDA device")
    cu = _abi(api)
    size = C.sizeof(api.REAL)
    out_dev = C.c_void_p()
    assert cu.grt_malloc(0, C.byref(out_dev), MAX_LEVELS*size) == 0
    rng = np.random.default_rng(12)
    rtol = RTOL if api.DTYPE == np.float64 else 2e-6
    try:
        # first call: 61 levels on a 0.1 cm-1 shortwave-sized grid (64 blocks per level);
        # second call: fewer levels, a short grid (fewer blocks per level than the first call
        # left behind in the scratch array), other values
        for levels, n, dw, scale in ((61, 499991, 0.1, 3.), (7, 1237, 0.01, 250.)):
            flux = np.ascontiguousarray(scale*rng.random((levels, n)), dtype=api.DTYPE)
            want, got = _both(api, cu, flux, dw, out_dev)
            exact = (0.5*(flux[:, :-1].astype(np.float64) + flux[:, 1:])*dw).sum(axis=1)
            np.testing.assert_allclose(got, want, rtol=rtol, atol=0.)
            np.testing.assert_allclose(got, exact, rtol=max(rtol, 1e-11), atol=0.)
            print(f"{levels} levels x {n} points: max rel difference to grt_integrate_flux "
                  f"{np.max(np.abs(got - want)/np.abs(want)):.2e}")
        # a single level of two points, and one point (no interval: the sum is zero)
        for n in (2, 1):
            flux = np.ascontiguousarray(rng.random((1, n)) + 1., dtype=api.DTYPE)
            want, got = _both(api, cu, flux, 0.5, out_dev)
            np.testing.assert_allclose(got, want, rtol=rtol, atol=0.)
        # beyond the static scratch: rejected with a message, nothing launched
        rc = cu.grt_integrate_flux_noalloc(0, MAX_LEVELS + 1, 100, 0.1, out_dev, out_dev)
        assert rc == 12                       # GRT_SM100_ERR
        buf = C.create_string_buffer(512)
        cu.grt_last_error(buf, 512)
        assert b"GRT_NOALLOC_MAX_LEVELS" in buf.value and b"exceed" in buf.value
        assert cu.grt_integrate_flux_noalloc(0, 3, 100, 0.1, None, out_dev) == 12
    finally:
        assert cu.grt_free(0, out_dev) == 0
