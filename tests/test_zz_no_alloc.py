"""dfu_split_cloud_static (include/dfu3d.h): the allocation-free variant of the modality split, whose scratch is a fixed
`static __device__` array inside the library.  It has no counterpart in the reference,
so the checker is the existing entry point dfu_split_cloud_sync on the same inputs: rows, order and counts
must be identical (the split copies rows: bit-exact, tolerance 0).  Called through the C ABI, twice in a row on different
inputs (the array keeps the previous call's contents), plus the size limit.  Collected last."""
import numpy as np
import pytest
import torch

from dfu3d_b200 import _lib

pytestmark = pytest.mark.gpu
RANGE = np.array([0, -40, -3, 70.4, 40, 1], np.float32)
MAX_ROWS = 4194304          # DFU_SPLIT_STATIC_MAX_ROWS


def cloud(seed, n, C=8):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-5, 80, size=(n, C)).astype(np.float32)
    c[:, 1] = rng.uniform(-45, 45, n)
    c[:, -1] = rng.choice([1.0, 2.0, 0.0], size=n, p=[0.7, 0.28, 0.02])
    c[5::11, 0] = np.float32(70.4)           # exactly on the inclusive range bounds
    c[6::13, 1] = np.float32(-40.0)
    return c


def split_sync(ctx, c, rg):
    n, C = c.shape
    o0 = torch.empty((max(n, 1), C), dtype=torch.float32, device=ctx.device)
    o1 = torch.empty_like(o0)
    cnt = np.zeros(2, np.int64)
    ctx.check(ctx.lib.dfu_split_cloud_sync(ctx.h, ctx.stream, _lib.dptr(c) if n else None, C, n,
                                           _lib.hptr(rg) if rg is not None else None, _lib.dptr(o0), _lib.dptr(o1), _lib.hptr(cnt)))
    return o0[:int(cnt[0])].cpu().numpy(), o1[:int(cnt[1])].cpu().numpy()


def split_static(ctx, c, rg, bufs, handle=True):
    """bufs = (out_lidar, out_pseudo, counts_dev) allocated by the caller beforehand: the call itself allocates nothing."""
    n, C = c.shape
    o0, o1, cnt = bufs
    rc = ctx.lib.dfu_split_cloud_static(ctx.h if handle else None, ctx.stream, _lib.dptr(c) if n else None, C, n,
                                        _lib.hptr(rg) if rg is not None else None, _lib.dptr(o0), _lib.dptr(o1), _lib.dptr(cnt))
    assert rc == 0, (rc, ctx.lib.dfu_last_error(ctx.h).decode())
    torch.cuda.synchronize(ctx.device)
    k = cnt.cpu().numpy()
    return o0[:int(k[0])].cpu().numpy(), o1[:int(k[1])].cpu().numpy()


def test_static_split_equals_sync_split_twice_in_a_row():
    ctx = _lib.default_context(0)
    C = 8
    # caller-owned outputs, sized once for the largest case; poisoned so that stale rows would show
    o0 = torch.full((300000, C), float('nan'), dtype=torch.float32, device=ctx.device)
    o1 = torch.full_like(o0, float('nan'))
    cnt = torch.full((2,), -7, dtype=torch.int64, device=ctx.device)
    cases = [(11, 300000, RANGE), (12, 4321, None), (13, 1, RANGE), (14, 0, None), (15, 257, RANGE)]
    clouds = [torch.from_numpy(cloud(s, n, C)).to(ctx.device) for s, n, _ in cases]
    torch.cuda.synchronize(ctx.device)
    for rep in range(2):                      # the whole sequence twice: big -> small -> big again over the same array
        for (s, n, rg), c in zip(cases, clouds):
            want = split_sync(ctx, c, rg)
            got = split_static(ctx, c, rg, (o0, o1, cnt), handle=(s % 2 == 1))
            assert got[0].shape == want[0].shape and got[1].shape == want[1].shape, (s, n)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (s, n)
    # back to back without anything in between: two different inputs through the shared array, then compare both
    a, b = clouds[0], clouds[1]
    oa0, oa1, ca = torch.empty_like(o0), torch.empty_like(o0), torch.empty_like(cnt)
    ob0, ob1, cb = torch.empty_like(o0), torch.empty_like(o0), torch.empty_like(cnt)
    torch.cuda.synchronize(ctx.device)
    free_mid = torch.cuda.mem_get_info(ctx.device)[0]
    for c, (x0, x1, k) in ((a, (oa0, oa1, ca)), (b, (ob0, ob1, cb))):
        rc = ctx.lib.dfu_split_cloud_static(ctx.h, ctx.stream, _lib.dptr(c), C, c.shape[0], _lib.hptr(RANGE),
                                            _lib.dptr(x0), _lib.dptr(x1), _lib.dptr(k))
        assert rc == 0
    torch.cuda.synchronize(ctx.device)
    assert torch.cuda.mem_get_info(ctx.device)[0] == free_mid           # the calls took no device memory
    for c, (x0, x1, k) in ((a, (oa0, oa1, ca)), (b, (ob0, ob1, cb))):
        want = split_sync(ctx, c, RANGE)
        kk = k.cpu().numpy()
        assert (int(kk[0]), int(kk[1])) == (want[0].shape[0], want[1].shape[0])
        assert np.array_equal(x0[:int(kk[0])].cpu().numpy(), want[0]) and np.array_equal(x1[:int(kk[1])].cpu().numpy(), want[1])


def test_static_split_rejects_sizes_beyond_the_fixed_scratch():
    ctx = _lib.default_context(0)
    tiny = torch.zeros((1, 8), dtype=torch.float32, device=ctx.device)
    cnt = torch.full((2,), -7, dtype=torch.int64, device=ctx.device)
    # the size check comes before anything is launched or read: the buffers are never touched
    rc = ctx.lib.dfu_split_cloud_static(ctx.h, ctx.stream, _lib.dptr(tiny), 8, MAX_ROWS + 1, None,
                                        _lib.dptr(tiny), _lib.dptr(tiny), _lib.dptr(cnt))
    assert rc == -3                                   # DFU_ERR_CAPACITY
    msg = ctx.lib.dfu_last_error(ctx.h).decode()
    assert 'DFU_SPLIT_STATIC_MAX_ROWS' in msg and str(MAX_ROWS) in msg
    torch.cuda.synchronize(ctx.device)
    assert cnt.cpu().tolist() == [-7, -7]
    assert ctx.lib.dfu_split_cloud_static(None, ctx.stream, _lib.dptr(tiny), 8, MAX_ROWS + 1, None,
                                          _lib.dptr(tiny), _lib.dptr(tiny), _lib.dptr(cnt)) == -3
    assert ctx.lib.dfu_split_cloud_static(ctx.h, ctx.stream, _lib.dptr(tiny), 8, 1, None,
                                          _lib.dptr(tiny), _lib.dptr(tiny), None) == -1      # DFU_ERR_ARG


def test_static_split_at_the_maximum_size():
    ctx = _lib.default_context(0)
    g = torch.Generator(device='cpu').manual_seed(5)
    c = torch.rand((MAX_ROWS, 4), generator=g, dtype=torch.float32) * 60
    c[:, 3] = torch.randint(0, 3, (MAX_ROWS,), generator=g).float()
    c = c.to(ctx.device)
    o0, o1 = torch.empty_like(c), torch.empty_like(c)
    cnt = torch.zeros(2, dtype=torch.int64, device=ctx.device)
    want = split_sync(ctx, c, RANGE)
    got = split_static(ctx, c, RANGE, (o0, o1, cnt))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert got[0].shape[0] + got[1].shape[0] <= MAX_ROWS
