"""Stage 3 parity through the C ABI: the voxel map vs the oracle's exact counts and octomap-rule pointer tree."""
import numpy as np
import pytest

from util import ramp_cloud, random_cloud

pytestmark = pytest.mark.gpu


def test_set_octree_like_reference_unit(ctx, pcs, O):
    # octomap_mesh_masking_unit.cpp:59-70: setOctree(cloud, 0.1 [,0,255,false]) on the grey ramp cloud
    cloud = ramp_cloud(pcs.POINT_DTYPE)
    vmap = ctx.create_map(0.1, 1024)
    vmap.integrate(cloud, None, 0, 255, False)
    keys, c0, c1, lo = vmap.export()
    tree = O.Octree(0.1)
    tree.insert_cloud(O.color_passthrough(cloud, 0, 255))
    tkeys, tvals = tree.leaf_voxels()
    assert keys.size > 0 and np.array_equal(np.sort(keys), tkeys)
    order = np.argsort(keys)
    assert np.array_equal(lo[order].view(np.uint32), tvals.view(np.uint32))  # log-odds: 1e-5 allowed, exact obtained
    assert (c0 == 0).all() and int(c1.sum()) == 255
    vmap.destroy()


@pytest.mark.parametrize("flt", [(0, 255, False), (50, 180, False), (50, 180, True)])
@pytest.mark.parametrize("res", [0.01, 0.005, 0.002, 0.1])
def test_integrate_random_cloud(ctx, pcs, O, flt, res):
    lower, upper, neg = flt
    rng = np.random.default_rng(int(res * 1e4) + lower)
    cloud = random_cloud(pcs.POINT_DTYPE, 40000, rng, extent=0.3)
    pose = O.synth_pose(17)
    vmap = ctx.create_map(res, 1 << 17)
    vmap.integrate(cloud, pose, lower, upper, neg)
    keys, c0, c1, lo = vmap.export()
    o_k, o_c0, o_c1 = O.count_voxels(cloud, res, pose, lower, upper, neg)
    assert np.array_equal(keys, o_k) and np.array_equal(c0, o_c0) and np.array_equal(c1, o_c1)
    vmap.destroy()


def test_heavy_collisions_and_clamp(ctx, pcs, O):
    """Thousands of points in a handful of voxels: warp-aggregated atomics must still count exactly; log-odds clamps."""
    rng = np.random.default_rng(5)
    cloud = random_cloud(pcs.POINT_DTYPE, 200000, rng, extent=0.011, nan_frac=0.0, inf_frac=0.0)
    vmap = ctx.create_map(0.01, 1024)
    vmap.integrate(cloud)
    keys, c0, c1, lo = vmap.export()
    o_k, o_c0, o_c1 = O.count_voxels(cloud, 0.01)
    assert keys.size <= 64 and np.array_equal(keys, o_k) and np.array_equal(c1, o_c1)
    assert int(c1.sum()) == 200000
    assert (lo.view(np.uint32) == 0x4060B4BA).all()  # clamp logodds(0.971)
    vmap.destroy()


def test_export_is_morton_ordered(ctx, pcs, O):
    rng = np.random.default_rng(6)
    cloud = random_cloud(pcs.POINT_DTYPE, 30000, rng, extent=2.0)
    vmap = ctx.create_map(0.05, 1 << 16)
    vmap.integrate(cloud)
    keys, _, _, _ = vmap.export()
    m = O.morton48(keys)
    assert (np.diff(m.astype(np.int64)) > 0).all()
    vmap.destroy()


def test_merge_is_linear(ctx, pcs, O):
    """map(A) merged into map(B) == map(A u B): the property the multi-GPU exchange relies on."""
    rng = np.random.default_rng(7)
    a = random_cloud(pcs.POINT_DTYPE, 20000, rng, extent=0.2)
    b = random_cloud(pcs.POINT_DTYPE, 30000, rng, extent=0.2)
    ma, mb, mab = (ctx.create_map(0.01, 1 << 16) for _ in range(3))
    ma.integrate(a, None, 0, 128)
    mb.integrate(b, None, 0, 128)
    mab.integrate(np.concatenate([a, b]), None, 0, 128)
    ka, a0, a1, _ = ma.export()
    mb.merge(ka, a0, a1)
    assert all(np.array_equal(x, y) for x, y in zip(mb.export(), mab.export()))
    for m in (ma, mb, mab):
        m.destroy()


def test_pack_by_owner_and_merge_records(ctx, pcs, O):
    rng = np.random.default_rng(8)
    cloud = random_cloud(pcs.POINT_DTYPE, 50000, rng, extent=0.5)
    src = ctx.create_map(0.01, 1 << 17)
    src.integrate(cloud, None, 0, 100)
    n_vox, _ = src.size()
    rec_dev = ctx.dev_alloc(n_vox * 16)
    for n_owners in (1, 2, 8):
        counts = src.pack_by_owner_dev(n_owners, rec_dev, n_vox)
        assert int(counts.sum()) == n_vox
        recs = np.empty(n_vox, pcs.capi.RECORD_DTYPE)
        ctx.d2h(recs, rec_dev)
        owners = np.array([pcs.capi.mix64(int(k)) % n_owners for k in recs["key48"]])
        bounds = np.concatenate([[0], np.cumsum(counts)]).astype(int)
        for o in range(n_owners):
            assert (owners[bounds[o]:bounds[o + 1]] == o).all()
        dst = ctx.create_map(0.01, 1 << 17)
        dst.merge_records_dev(rec_dev, n_vox)
        assert all(np.array_equal(x, y) for x, y in zip(dst.export(), src.export()))
        dst.destroy()
    ctx.dev_free(rec_dev)
    src.destroy()


def test_capacity_overflow_is_reported(ctx, pcs):
    rng = np.random.default_rng(9)
    cloud = random_cloud(pcs.POINT_DTYPE, 20000, rng, extent=5.0, nan_frac=0, inf_frac=0)
    vmap = ctx.create_map(0.01, 256)  # 1024 slots, 20000 distinct voxels
    vmap.integrate(cloud)
    with pytest.raises(pcs.PcsError) as e:
        vmap.size()
    assert e.value.status == pcs.capi.PCS_ERR_CAPACITY
    vmap.destroy()


def test_empty_map_and_bad_args(ctx, pcs):
    vmap = ctx.create_map(0.01, 16)
    assert vmap.size() == (0, 0)
    k, c0, c1, lo = vmap.export()
    assert k.size == 0
    vmap.integrate(np.zeros(0, pcs.POINT_DTYPE))
    with pytest.raises(pcs.PcsError):
        vmap.integrate(ramp_cloud(pcs.POINT_DTYPE), None, 10, 5)
    with pytest.raises(pcs.PcsError):
        ctx.create_map(0.0, 16)
    vmap.destroy()


def test_push_to_owners_partitions_the_map(ctx, pcs, O):
    """Peer-memory exchange kernel with the owner tables in this process: the owners' slices partition the local map by
    mix64(key48) % n_owners and their union equals it (the multi-GPU run maps the same tables through CUDA IPC)."""
    rng = np.random.default_rng(21)
    cloud = random_cloud(pcs.POINT_DTYPE, 60000, rng, extent=0.6)
    local = ctx.create_map(0.01, 1 << 17)
    local.integrate(cloud, None, 0, 120)
    want = local.export()
    for n_owners in (1, 3, 8):
        owners = [ctx.create_map(0.01, 1 << 17) for _ in range(n_owners)]
        local.push_to_owners_dev(owners)
        local.push_to_owners_dev(owners)  # pushing twice doubles every counter: adds, not overwrites
        total = ctx.create_map(0.01, 1 << 17)
        n_seen = 0
        for o, m in enumerate(owners):
            k, c0, c1, _ = m.export()
            assert all(pcs.capi.mix64(int(x)) % n_owners == o for x in k[:: max(1, k.size // 200)])
            n_seen += k.size
            total.merge(k, c0, c1)
        assert n_seen == want[0].size
        k, c0, c1, _ = total.export()
        assert np.array_equal(k, want[0]) and np.array_equal(c0, 2 * want[1]) and np.array_equal(c1, 2 * want[2])
        for m in owners + [total]:
            m.destroy()
    with pytest.raises(pcs.PcsError):
        local.push_to_owners_dev([local])
    # clear_local: the push also leaves the local map empty and ready for the next scan
    a, b = ctx.create_map(0.01, 1 << 17), ctx.create_map(0.01, 1 << 17)
    want = local.export()
    local.push_to_owners_dev([a], clear_local=True)
    assert local.size() == (0, 0)
    assert all(np.array_equal(x, y) for x, y in zip(a.export(), want))
    local.integrate(cloud, None, 0, 120)  # the emptied table takes the same scan again
    assert all(np.array_equal(x, y) for x, y in zip(local.export(), want))
    local.push_to_owners_dev([b], clear_local=True)
    assert all(np.array_equal(x, y) for x, y in zip(b.export(), want))
    a.destroy()
    b.destroy()
    local.destroy()


def test_map_stats_counters(ctx, pcs, O):
    """pcs_map_get_stats: voxels touched / occupied, rejected and labelled point counts, load factor."""
    rng = np.random.default_rng(31)
    cloud = random_cloud(pcs.POINT_DTYPE, 50000, rng, extent=0.5)
    vmap = ctx.create_map(0.02, 1 << 16)
    vmap.integrate(cloud, None, 0, 128)
    keys, c0, c1, _ = vmap.export()
    st = vmap.stats()
    assert st["voxels"] == keys.size and st["occupied"] == int((c1 > 0).sum())
    assert st["points_rejected"] == int(c0.sum()) and st["points_labelled"] == int(c1.sum())
    assert st["table_slots"] == 1 << 17 and abs(st["load_factor"] - keys.size / (1 << 17)) < 1e-12
    o_k, o_c0, o_c1 = O.count_voxels(cloud, 0.02, None, 0, 128)
    assert st["points_labelled"] == int(o_c1.sum()) and st["points_rejected"] == int(o_c0.sum())
    vmap.clear()
    assert vmap.stats()["voxels"] == 0 and vmap.stats()["points_labelled"] == 0
    vmap.destroy()
