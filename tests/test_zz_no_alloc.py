"""rnb_raymarching_train_noalloc (include/radnerf_b200.h): the training marcher with its per-ray scratch in a static
device array inside the library, for callers that cannot call cudaMalloc.  No counterpart in the reference;
the expectation is the EXISTING two-phase entry pair rnb_raymarching_train_count +
rnb_raymarching_train_write on the same inputs.  Integer outputs (rays_a, counter) and the sample payload are compared
bit-exactly -- the same kernels run, only the scratch differs.  Two calls in a row with different inputs: the array keeps
the first call's contents and the second call must not see them.  Collected last (file name), so that a failure here cannot
hide another test under `pytest -x`; passes alone in a fresh process and after the other tests.
"""
import ctypes

import numpy as np
import pytest
import torch

from conftest import CONFIGS, get_scene, make_rays
from radnerf_b200 import _lib
from radnerf_b200._lib import call, f32, i32, i64, stream_arg
from radnerf_b200 import synthetic as syn

pytestmark = pytest.mark.gpu
MAX_SAMPLES = 1024
MAX_FLOATS = 8 * 1024 * 1024      # RNB_NOALLOC_MARCH_MAX_FLOATS


def _inputs(scale, n, seed):
    sc, bf = get_scene(scale)
    o, d, h = make_rays(scale, n, seed=seed)
    noise = np.random.RandomState(seed + 1).rand(n).astype(np.float32)
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    return cu(o), cu(d), cu(h), cu(bf), cu(noise)


def _two_phase(o, d, h, bf, noise, cascades, scale, esf):
    """the existing entry points, with a caller-allocated t_scratch"""
    n, dev = o.shape[0], o.device
    rays_a = torch.empty(n, 3, dtype=torch.int64, device=dev)
    counter = torch.empty(2, dtype=torch.int32, device=dev)
    t_scratch = torch.empty(n * MAX_SAMPLES, dtype=torch.float32, device=dev)
    call("rnb_raymarching_train_count", o, d, h, bf, i32(cascades), f32(scale), f32(esf), noise, i32(128), i32(MAX_SAMPLES),
         i64(n), rays_a, counter, t_scratch, stream_arg(dev))
    total = int(counter[0])
    out = [torch.empty(total, 3, device=dev), torch.empty(total, 3, device=dev), torch.empty(total, device=dev), torch.empty(total, device=dev)]
    call("rnb_raymarching_train_write", o, d, f32(scale), f32(esf), i32(128), i32(MAX_SAMPLES), i64(n), rays_a, t_scratch,
         i64(total), *out, stream_arg(dev))
    return rays_a, counter, out


def _no_alloc(o, d, h, bf, noise, cascades, scale, esf, cap):
    n, dev = o.shape[0], o.device
    rays_a = torch.empty(n, 3, dtype=torch.int64, device=dev)
    counter = torch.empty(2, dtype=torch.int32, device=dev)
    out = [torch.full((cap, 3), -7.0, device=dev), torch.full((cap, 3), -7.0, device=dev), torch.full((cap,), -7.0, device=dev),
           torch.full((cap,), -7.0, device=dev)]
    call("rnb_raymarching_train_noalloc", o, d, h, bf, i32(cascades), f32(scale), f32(esf), noise, i32(128), i32(MAX_SAMPLES),
         i64(n), rays_a, counter, i64(cap), *out, stream_arg(dev))
    return rays_a, counter, out


def _same_bits(a, b):
    return torch.equal(a.contiguous().view(torch.int32), b.contiguous().view(torch.int32))


@pytest.mark.parametrize("cfg", ["bounded", "unbounded"])
def test_no_alloc_marcher_matches_two_phase_on_two_calls_in_a_row(cfg):
    scale, esf = CONFIGS[cfg]
    cascades = syn.cascades_for_scale(scale)
    # second call: fewer rays and other rays / noise, so rows of the shared scratch hold stale entries of the first call
    for n, seed in ((8192, 42), (5000, 77)):
        args = _inputs(scale, n, seed)
        ra, ca, pa = _two_phase(*args, cascades, scale, esf)
        total = int(ca[0])
        assert total > 0
        rb, cb, pb = _no_alloc(*args, cascades, scale, esf, cap=total + 1000)
        torch.cuda.synchronize()
        assert torch.equal(ra, rb) and torch.equal(ca, cb)                       # counts, offsets, indices: bit-exact
        for x, y in zip(pa, pb):
            assert _same_bits(x, y[:total])                                      # xyzs, dirs, deltas, ts: bit-exact
            assert bool((y[total:] == -7.0).all())                               # nothing written past the samples


def test_no_alloc_marcher_drops_samples_beyond_cap_and_rejects_oversize():
    scale, esf = CONFIGS["bounded"]
    args = _inputs(scale, 2048, 5)
    ra, ca, pa = _two_phase(*args, 1, scale, esf)
    total = int(ca[0])
    cap = total // 2
    rb, cb, pb = _no_alloc(*args, 1, scale, esf, cap=cap)
    torch.cuda.synchronize()
    assert torch.equal(ra, rb) and int(cb[0]) == total                           # the true total is still reported
    for x, y in zip(pa, pb):
        assert _same_bits(x[:cap], y)
    # beyond the built-in scratch: a clear error, nothing launched
    n_big = MAX_FLOATS // MAX_SAMPLES + 1
    o = torch.zeros(n_big, 3, device="cuda"); h = torch.zeros(n_big, 2, device="cuda"); nz = torch.zeros(n_big, device="cuda")
    rays_a = torch.empty(n_big, 3, dtype=torch.int64, device="cuda"); counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    one = torch.empty(3, device="cuda")
    with pytest.raises(_lib.RadNerfB200Error, match="exceeds the built-in scratch"):
        call("rnb_raymarching_train_noalloc", o, o, h, args[3], i32(1), f32(scale), f32(esf), nz, i32(128), i32(MAX_SAMPLES),
             i64(n_big), rays_a, counter, i64(1), one, one, one, one, stream_arg(o.device))
    # exactly at the limit is accepted
    n_max = MAX_FLOATS // MAX_SAMPLES
    call("rnb_raymarching_train_noalloc", o[:n_max], o[:n_max], h[:n_max], args[3], i32(1), f32(scale), f32(esf), nz[:n_max], i32(128),
         i32(MAX_SAMPLES), i64(n_max), rays_a, counter, i64(0), None, None, None, None, stream_arg(o.device))
    torch.cuda.synchronize()
    assert int(counter[1]) == n_max


def test_no_alloc_entry_allocates_nothing():
    """the driver's free-memory figure does not move across calls (after the first, which may load the module lazily)"""
    scale, esf = CONFIGS["bounded"]
    args = _inputs(scale, 4096, 9)
    _, c, _ = _two_phase(*args, 1, scale, esf)
    cap = int(c[0])
    n, dev = 4096, args[0].device
    rays_a = torch.empty(n, 3, dtype=torch.int64, device=dev); counter = torch.empty(2, dtype=torch.int32, device=dev)
    out = [torch.empty(cap, 3, device=dev), torch.empty(cap, 3, device=dev), torch.empty(cap, device=dev), torch.empty(cap, device=dev)]

    def run():
        call("rnb_raymarching_train_noalloc", *args[:4], i32(1), f32(scale), f32(esf), args[4], i32(128), i32(MAX_SAMPLES), i64(n),
             rays_a, counter, i64(cap), *out, stream_arg(dev))
    run(); torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(5):
        run()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0
