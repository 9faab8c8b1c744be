"""pcs_color_passthrough_noalloc_dev: the label filter for hosts that cannot let the library allocate (a per-frame
callback that must not stall the device on cudaMalloc, or a GPU whose memory another process has already claimed).

The caller owns every buffer; the entry point keeps its block counts in a fixed array inside the library.  The test
calls it through the C ABI twice in a row -- the second call starts with the first call's counts still in that
array -- and compares each result, byte for byte like the rest of the suite, with pcs_color_passthrough.
Collected last (file name) so that a -x run has already executed every other test if this one fails."""
import numpy as np
import pytest

from util import random_cloud, same_bytes

pytestmark = pytest.mark.gpu


def run_noalloc(ctx, pcs, cloud, lower, upper, neg, bufs):
    d_in, d_out, d_cnt = bufs
    n = cloud.size
    ctx.h2d(d_in, cloud)
    ctx.color_passthrough_noalloc_dev(d_in, n, lower, upper, neg, d_out, d_cnt)
    cnt = np.zeros(1, np.uint64)
    ctx.d2h(cnt, d_cnt)  # d2h synchronises the context's stream
    out = np.empty(int(cnt[0]), pcs.POINT_DTYPE)
    if out.size:
        ctx.d2h(out, d_out)
    return out


def test_two_calls_in_a_row_match_the_allocating_entry_point(ctx, pcs, O):
    rng = np.random.default_rng(2023)
    n1, n2 = 300007, 4099  # big first, small second: stale counts of call 1 sit beyond call 2's blocks
    a = random_cloud(pcs.POINT_DTYPE, n1, rng)
    b = random_cloud(pcs.POINT_DTYPE, n2, rng)
    a["pad"][:, 0] = np.arange(n1)
    b["pad"][:, 1] = np.arange(n2)
    bufs = (ctx.dev_alloc(n1 * 32), ctx.dev_alloc(n1 * 32), ctx.dev_alloc(8))  # caller-owned device memory
    launches0 = ctx.stats()["kernel_launches"]
    got_a = run_noalloc(ctx, pcs, a, 30, 200, False, bufs)
    got_b = run_noalloc(ctx, pcs, b, 90, 160, True, bufs)
    assert ctx.stats()["kernel_launches"] - launches0 == 6  # count, scan, scatter per call: the CUDA path ran
    want_a = ctx.color_passthrough(a, 30, 200, False)
    want_b = ctx.color_passthrough(b, 90, 160, True)
    assert 0 < want_a.size < n1 and 0 < want_b.size < n2
    assert got_a.tobytes() == want_a.tobytes()
    assert got_b.tobytes() == want_b.tobytes()
    # and against the CPU oracle directly, so the chain does not rest on the sibling entry point alone
    assert same_bytes(got_a, O.color_passthrough(a, 30, 200, False))
    assert same_bytes(got_b, O.color_passthrough(b, 90, 160, True))
    # empty input: the count is cleared, nothing else is touched
    ctx.color_passthrough_noalloc_dev(bufs[0], 0, 0, 255, False, bufs[1], bufs[2])
    cnt = np.ones(1, np.uint64)
    ctx.d2h(cnt, bufs[2])
    assert cnt[0] == 0
    for p in bufs:
        ctx.dev_free(p)


def test_sizes_beyond_the_fixed_array_are_refused(ctx, pcs):
    bufs = (ctx.dev_alloc(64), ctx.dev_alloc(64), ctx.dev_alloc(8))
    with pytest.raises(pcs.PcsError) as e:  # rejected before any memory is touched
        ctx.color_passthrough_noalloc_dev(bufs[0], pcs.capi.NOALLOC_MAX_POINTS + 1, 0, 255, False, bufs[1], bufs[2])
    assert e.value.status == pcs.capi.PCS_ERR_CAPACITY and "PCS_NOALLOC_MAX_POINTS" in str(e.value)
    with pytest.raises(pcs.PcsError) as e:
        ctx.color_passthrough_noalloc_dev(bufs[0], 2, 9, 3, False, bufs[1], bufs[2])
    assert e.value.status == pcs.capi.PCS_ERR_BAD_ARG
    with pytest.raises(pcs.PcsError) as e:  # in-place filtering would corrupt the output silently: refused
        ctx.color_passthrough_noalloc_dev(bufs[0], 2, 0, 255, False, bufs[0] + 32, bufs[2])
    assert e.value.status == pcs.capi.PCS_ERR_BAD_ARG and "overlap" in str(e.value)
    for p in bufs:
        ctx.dev_free(p)
