"""svo_gpu_bucket_static: the bucketing entry point for callers that must not allocate device memory.

Who needs it: a tracking loop that has promised its scheduler not to allocate once frames flow (cudaMalloc / cudaFree take
driver-wide locks and may wait for the whole device), and plug-ins living inside an application whose own pool already owns
all of the GPU's memory, where a cudaMalloc from a guest library simply fails.  Those callers cannot use svo_gpu_create (it
allocates every context buffer) nor svo_gpu_bucket (it grows its bucket tables on demand).  They hand in device buffers they
already own, and the per-bucket tables sit in a fixed 1 MiB device array compiled into libsvo_gpu.so.

What is checked, through the C ABI: on the same inputs the entry point returns exactly what svo_gpu_bucket returns (points,
ages, order, duplicates -- integer / copy work, so the tolerance is zero as in every other bucketing test), on calls made
back to back so that each one finds the previous call's tables in the array: a large table followed by a small one, a
repeated call, and a call after one that overflowed its output capacity.  Also that a grid too fine for the static array, a
point outside the table and host pointers are refused with an error code and a message rather than computed wrongly.

(The test itself allocates its device buffers with torch -- the test is not the caller that cannot.)  This file is named so
that pytest collects it last: `pytest -x` stops at the first failure, and a failure here must not hide the parity tests.
"""
import numpy as np
import pytest

from svo_b200 import capi
from test_gpu_parity import _bucket_cases

pytestmark = pytest.mark.gpu


def _run_static(xy, ages, w, h, bs, k, capacity=None, stream=0):
    import torch
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    ages = np.ascontiguousarray(ages, np.int32).reshape(-1)
    n = len(xy)
    if capacity is None:
        capacity = max(n, (h // bs + 1) * (w // bs + 1) * max(k, 1))
    dev = torch.device("cuda", 0)
    d_xy = torch.from_numpy(xy).to(dev) if n else torch.zeros((1, 2), dtype=torch.float32, device=dev)
    d_ag = torch.from_numpy(ages).to(dev) if n else torch.zeros(1, dtype=torch.int32, device=dev)
    o_xy = torch.full((max(capacity, 1), 2), -7.0, dtype=torch.float32, device=dev)
    o_ag = torch.full((max(capacity, 1),), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    m = capi.bucket_static(d_xy.data_ptr(), d_ag.data_ptr(), n, w, h, bs, k, o_xy.data_ptr(), o_ag.data_ptr(), capacity,
                           stream)
    # the call has returned: results are valid without any further synchronisation on our side
    out_xy, out_ag = o_xy.cpu().numpy(), o_ag.cpu().numpy()
    assert (out_xy[m:] == -7.0).all() and (out_ag[m:] == -7).all(), "wrote past the reported count"
    assert np.array_equal(d_xy.cpu().numpy()[:n], xy) and np.array_equal(d_ag.cpu().numpy()[:n], ages), "inputs modified"
    return out_xy[:m].copy(), out_ag[:m].copy()


@pytest.fixture(scope="module")
def ctx():
    c = capi.Context(device=0, max_w=1280, max_h=512, max_levels=4, max_points=40000, max_batch=1)
    yield c
    c.close()


def test_static_matches_context_entry_point_back_to_back(ctx):
    """Every bucketing case of the parity suite, twice in a row each and in an order that alternates table sizes, against
    svo_gpu_bucket: identical output each time, whatever the previous call left in the static array."""
    cases = list(_bucket_cases())
    order = cases + cases[::-1]          # ... dense, dense, integer, ...: includes the same case on consecutive calls
    for name, xy, ages, w, h, bs, k in order:
        want_xy, want_ag = ctx.bucket(xy, ages, w, h, bs, k)
        for rep in range(2):
            got_xy, got_ag = _run_static(xy, ages, w, h, bs, k)
            assert np.array_equal(got_xy, want_xy) and np.array_equal(got_ag, want_ag), (name, rep)


def test_static_on_a_side_stream_and_without_ages(ctx):
    import torch
    rng = np.random.default_rng(3)
    xy = rng.uniform([0, 0], [1241, 376], size=(7000, 2)).astype(np.float32)
    want_xy, want_ag = ctx.bucket(xy, np.zeros(len(xy), np.int32), 1241, 376, 32, 5)
    st = torch.cuda.Stream()
    d_xy = torch.from_numpy(xy).cuda()
    cap = len(xy)
    o_xy = torch.zeros((cap, 2), dtype=torch.float32, device="cuda")
    o_ag = torch.full((cap,), 9, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):
        m = capi.bucket_static(d_xy.data_ptr(), None, len(xy), 1241, 376, 32, 5, o_xy.data_ptr(), o_ag.data_ptr(), cap,
                               st.cuda_stream)
        assert np.array_equal(o_xy.cpu().numpy()[:m], want_xy) and np.array_equal(o_ag.cpu().numpy()[:m], want_ag)


def test_static_capacity_overflow_then_clean_call(ctx):
    rng = np.random.default_rng(4)
    xy = rng.uniform([0, 0], [640, 480], size=(4000, 2)).astype(np.float32)
    ages = rng.integers(0, 30, 4000).astype(np.int32)
    want_xy, want_ag = ctx.bucket(xy, ages, 640, 480, 40, 4)
    assert len(want_xy) > 100
    with pytest.raises(capi.SvoGpuError) as e:
        _run_static(xy, ages, 640, 480, 40, 4, capacity=100)
    assert e.value.code == capi.E_CAPACITY and e.value.required == len(want_xy)
    got_xy, got_ag = _run_static(xy, ages, 640, 480, 40, 4)
    assert np.array_equal(got_xy, want_xy) and np.array_equal(got_ag, want_ag)


def test_static_refuses_what_it_cannot_do(ctx):
    import torch
    one = np.array([[5.0, 5.0]], np.float32)
    # a grid whose tables exceed the static array: 4096 x 4096 with 4-px buckets of 8 = 1025 * 1025 * 10 ints = 42 MB
    with pytest.raises(capi.SvoGpuError) as e:
        _run_static(one, np.zeros(1, np.int32), 4096, 4096, 4, 8, capacity=16)
    assert e.value.code == capi.E_CAPACITY and "static" in str(e.value)
    # the largest grid that does fit still works right after the refusal (3840 x 2160, 16-px buckets of 5)
    rng = np.random.default_rng(5)
    xy = rng.uniform([0, 0], [3840, 2160], size=(30000, 2)).astype(np.float32)
    ages = rng.integers(0, 9, 30000).astype(np.int32)
    with capi.Context(device=0, max_w=3840, max_h=2160, max_levels=1, max_points=200000, max_batch=1) as big:
        want_xy, want_ag = big.bucket(xy, ages, 3840, 2160, 16, 5)
    got_xy, got_ag = _run_static(xy, ages, 3840, 2160, 16, 5)
    assert np.array_equal(got_xy, want_xy) and np.array_equal(got_ag, want_ag)
    # a point outside the bucket table: the same refusal as svo_gpu_bucket (undefined behaviour in the reference)
    with pytest.raises(capi.SvoGpuError) as e:
        _run_static(np.array([[-500.0, 10.0]], np.float32), np.zeros(1, np.int32), 640, 480, 50, 4)
    assert e.value.code == capi.E_ARG
    # host memory where device memory is required
    out_xy = torch.zeros((4, 2), dtype=torch.float32, device="cuda")
    out_ag = torch.zeros(4, dtype=torch.int32, device="cuda")
    with pytest.raises(capi.SvoGpuError) as e:
        capi.bucket_static(one.ctypes.data, None, 1, 640, 480, 50, 4, out_xy.data_ptr(), out_ag.data_ptr(), 4)
    assert e.value.code == capi.E_ARG and "device memory" in str(e.value)
    # and the array is still usable afterwards
    got_xy, got_ag = _run_static(one, np.array([3], np.int32), 640, 480, 50, 4)
    assert np.array_equal(got_xy, one) and np.array_equal(got_ag, np.array([3], np.int32))
