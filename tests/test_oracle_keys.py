"""What can be pinned about the octomap arithmetic without octomap: the voxel key formula against exact
rational arithmetic on every voxel boundary of three resolutions, and the log-odds constants against their definition."""
import math

import numpy as np
import pytest

from util import expected_key_histogram, key_boundary_cloud


def test_logodds_constants_follow_from_double_log_then_cast(O):
    """octomap: logodds(p) = (float) log(p / (1 - p)) -- a double log, cast once.  The float sequence the map reports
    (0x3f58e883 ... 0x4060b4ba) follows from that and from float accumulation with a clamp; the all-float variant (logf)
    gives different bits, so the distinction is observable."""
    hit, clamp = np.float32(math.log(0.7 / 0.3)), np.float32(math.log(0.971 / 0.029))
    assert hit.view(np.uint32) == 0x3F58E883 and clamp.view(np.uint32) == 0x4060B4BA
    v, seq = np.float32(0), []
    for _ in range(7):
        if v < clamp:
            v = np.float32(v + hit)
            v = clamp if v > clamp else v
        seq.append(int(v.view(np.uint32)))
    assert seq == [0x3F58E883, 0x3FD8E883, 0x4022AE62, 0x4058E883, 0x4060B4BA, 0x4060B4BA, 0x4060B4BA]
    assert [int(np.float32(O.logodds_after_hits(c)).view(np.uint32)) for c in range(1, 8)] == seq
    f = np.float32
    hit_f = f(np.log(f(f(0.7) / (f(1) - f(0.7)))))
    clamp_f = f(np.log(f(f(0.971) / (f(1) - f(0.971)))))
    assert hit_f.view(np.uint32) == 0x3F58E882 and clamp_f.view(np.uint32) == 0x4060B4BC  # not what octomap computes


@pytest.mark.parametrize("res", [0.01, 0.005, 0.002])
def test_every_voxel_boundary_against_exact_arithmetic(O, res):
    """Coordinates at k * res (as float32, and one float32 ulp either side) for all 65536 keys of an axis, plus values just
    outside the key range: the key must be floor(fl64(1 / res) * fl64(x)) + 32768 with the product rounded once to double.
    Expected keys come from fractions.Fraction; the oracle's histogram of keys must equal the expected one."""
    cloud = key_boundary_cloud(O.POINT_DTYPE, res)
    want_keys, want_counts = expected_key_histogram(cloud["x"], res)
    keys, c0, c1 = O.count_voxels(cloud, res)
    got = dict(zip((keys & np.uint64(0xFFFF)).tolist(), (c0.astype(np.uint64) + c1).tolist()))
    assert ((keys >> np.uint64(16)) == np.uint64((32768 << 16) | 32768)).all()  # y = z = 0 -> key 32768
    assert got == dict(zip(want_keys.tolist(), want_counts.tolist()))
    assert len(got) == 65536 and sum(got.values()) < cloud.size  # every key hit; the out-of-range probes were culled
