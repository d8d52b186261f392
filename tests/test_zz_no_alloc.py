"""zkb_bn_train_sign_static: the BatchNorm(train)+sign entry point for callers that cannot
allocate device memory (operator callbacks inside another runtime that owns the allocator, loops
that must not hit the device-wide synchronisation of cudaMalloc/cudaFree).  Its statistics live in
a fixed array inside the library, shared by all calls on a GPU, so the test issues two calls in a
row with different data and checks each against the ordinary three-call sequence
(zkb_bn_stats -> zkb_bn_finalize -> zkb_bn_apply_sign) with caller-provided buffers.

Collected last on purpose (file name): a failure here cannot hide the rest of the suite under -x.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

EPS, MOM = 1e-3, 0.99


def _reference(zn, x, is_i16, C, beta, bessel):
    rows = x.numel() // C
    dev = x.device
    acc = torch.zeros(2, C, dtype=torch.float64, device=dev)
    mean = torch.empty(C, dtype=torch.float32, device=dev)
    rstd = torch.empty(C, dtype=torch.float32, device=dev)
    mm = torch.zeros(C, dtype=torch.float32, device=dev)
    mv = torch.ones(C, dtype=torch.float32, device=dev)
    s = torch.empty(rows, C, dtype=torch.bfloat16, device=dev)
    a = torch.empty(rows, C, dtype=torch.float32, device=dev)
    zn.call("zkb_bn_stats", zn.ptr(x), is_i16, rows, C, acc[0].data_ptr(), acc[1].data_ptr(), 0)
    zn.call("zkb_bn_finalize", acc[0].data_ptr(), acc[1].data_ptr(), float(rows), EPS, MOM, bessel, C,
            zn.ptr(mean), zn.ptr(rstd), zn.ptr(mm), zn.ptr(mv), 0)
    zn.call("zkb_bn_apply_sign", zn.ptr(x), is_i16, rows, C, zn.ptr(mean), zn.ptr(rstd), zn.ptr(beta),
            zn.ptr(s), zn.ptr(a), 0, 0)
    torch.cuda.synchronize()
    return s, a, mm, mv


def _static(zn, x, is_i16, C, beta, bessel):
    rows = x.numel() // C
    dev = x.device
    mm = torch.zeros(C, dtype=torch.float32, device=dev)
    mv = torch.ones(C, dtype=torch.float32, device=dev)
    s = torch.empty(rows, C, dtype=torch.bfloat16, device=dev)
    a = torch.empty(rows, C, dtype=torch.float32, device=dev)
    zn.call("zkb_bn_train_sign_static", zn.ptr(x), is_i16, rows, C, zn.ptr(beta), EPS, MOM, bessel,
            zn.ptr(mm), zn.ptr(mv), zn.ptr(s), zn.ptr(a), 0)
    torch.cuda.synchronize()
    return s, a, mm, mv


def test_static_scratch_matches_three_call_sequence_twice(native_lib):
    from zookeeper_b200 import native as zn

    g = torch.Generator(device="cpu").manual_seed(3)
    # call 1: integer conv output, 512 channels; call 2: float map, fewer channels and other values,
    # so stale accumulators from call 1 would show up in call 2
    x1 = torch.randint(-1152, 1153, (5000, 512), generator=g, dtype=torch.int16).cuda()
    b1 = torch.randn(512, generator=g).cuda()
    x2 = (torch.randn(3001, 128, generator=g) * 3 + 0.5).cuda()
    b2 = torch.randn(128, generator=g).cuda()
    for x, is_i16, C, beta, bessel in ((x1, 1, 512, b1, 1), (x2, 0, 128, b2, 0)):
        ref = _reference(zn, x, is_i16, C, beta, bessel)
        got = _static(zn, x, is_i16, C, beta, bessel)
        assert torch.equal(got[0].view(torch.int16), ref[0].view(torch.int16)), "sign output differs"
        # same kernels, same summation order apart from atomics: tolerance of the suite (1e-4 rel)
        for gt, rf, what in zip(got[1:], ref[1:], ("normalised map", "moving mean", "moving var")):
            err = (gt.double() - rf.double()).norm() / rf.double().norm().clamp_min(1e-30)
            assert err.item() < 1e-4, f"{what}: rel err {err.item():.3g}"


def _cmax(zn, is_i16):
    lib = zn.load()
    lib.zkb_bn_train_sign_static_max_channels.restype = ctypes.c_int
    lib.zkb_bn_train_sign_static_max_channels.argtypes = [ctypes.c_int]
    return lib.zkb_bn_train_sign_static_max_channels(is_i16)


@pytest.mark.parametrize("is_i16", [1, 0])
def test_static_scratch_at_exactly_the_documented_limit(native_lib, is_i16):
    """The widest layer the entry point accepts (2048 channels of int16 input, 1024 of f32): same
    results as the three-call sequence, so the whole array is usable as documented."""
    from zookeeper_b200 import native as zn

    C = _cmax(zn, is_i16)
    assert C == (2048 if is_i16 else 1024)
    g = torch.Generator(device="cpu").manual_seed(17 + is_i16)
    if is_i16:
        x = torch.randint(-4608, 4609, (777, C), generator=g, dtype=torch.int16).cuda()
    else:
        x = (torch.randn(777, C, generator=g) * 5 - 1).cuda()
    beta = torch.randn(C, generator=g).cuda()
    ref = _reference(zn, x, is_i16, C, beta, 1)
    got = _static(zn, x, is_i16, C, beta, 1)
    assert torch.equal(got[0].view(torch.int16), ref[0].view(torch.int16))
    for gt, rf in zip(got[1:], ref[1:]):
        err = (gt.double() - rf.double()).norm() / rf.double().norm().clamp_min(1e-30)
        assert err.item() < 1e-4


@pytest.mark.parametrize("is_i16", [1, 0])
def test_static_scratch_rejects_what_it_cannot_hold(native_lib, is_i16):
    """One channel vector beyond the limit, twice the limit, and a width that is not a power-of-two
    number of vectors: refused before any launch, with a message that names the rule."""
    from zookeeper_b200 import native as zn

    cmax = _cmax(zn, is_i16)
    vec = 8 if is_i16 else 4
    dt = torch.int16 if is_i16 else torch.float32
    for C, msg in ((cmax + vec, "exceeds the built-in scratch"), (2 * cmax, "exceeds the built-in scratch"),
                   (3 * vec, "times a power of two")):
        x = torch.zeros(4, C, dtype=dt, device="cuda")
        beta = torch.zeros(C, dtype=torch.float32, device="cuda")
        s = torch.empty(4, C, dtype=torch.bfloat16, device="cuda")
        with pytest.raises(zn.ZkbError, match=msg):
            zn.call("zkb_bn_train_sign_static", zn.ptr(x), is_i16, 4, C, zn.ptr(beta), EPS, MOM, 0, 0, 0, zn.ptr(s), 0, 0)
