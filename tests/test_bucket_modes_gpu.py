"""Bucketing has two device formulations of FeatureSet::bucketingFeatures (reference src/bucket.cpp:24-111): three
global-memory kernels (clear, insert, emit) and one CTA per lane with the lane's bucket table in shared memory (used when
several lanes run at once).  SVO_GPU_BUCKET_MODE (read by svo_gpu_create) forces either; both must reproduce the golden
fixtures of the reference's own bucket.cpp and the oracle, bit for bit -- points, ages, order, duplicates, error behaviour."""
import os

import numpy as np
import pytest

from svo_b200 import capi
from test_gpu_parity import _bucket_cases

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(params=["1", "2"], ids=["global_kernels", "lane_kernel"])
def mctx(request, monkeypatch):
    monkeypatch.setenv("SVO_GPU_BUCKET_MODE", request.param)
    c = capi.Context(device=0, max_w=1280, max_h=512, max_levels=4, max_points=40000, max_batch=1)
    yield c
    c.close()


def test_bucket_modes_golden(mctx):
    g = np.load(os.path.join(GOLD, "bucket.npz"))
    for name in g["cases"]:
        w, h, bs, k = (int(v) for v in g["cfg_" + name])
        gx, ga = mctx.bucket(g["in_xy_" + name], g["in_age_" + name], w, h, bs, k)
        assert np.array_equal(gx, g["out_xy_" + name]) and np.array_equal(ga, g["out_age_" + name]), name


def test_bucket_modes_vs_oracle(mctx, oracle):
    for name, xy, ages, w, h, bs, k in _bucket_cases():
        gx, ga = mctx.bucket(xy, ages, w, h, bs, k)
        ox, oa = oracle.orc_bucket(xy, ages, w, h, bs, k)
        assert np.array_equal(gx, ox) and np.array_equal(ga, oa), name


def test_bucket_modes_errors(mctx):
    with pytest.raises(capi.SvoGpuError) as e:
        mctx.bucket(np.array([[-500.0, 10.0]], np.float32), np.zeros(1, np.int32), 640, 480, 50, 4)
    assert e.value.code == capi.E_ARG
