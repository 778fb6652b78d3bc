"""mtd_waypoints_paths_noalloc (include/mazetd_b200.h): the allocation-free variant of the batched Waypoints::update +
path walks, whose scratch (layers the walker warps chase, map counter) is a fixed array inside the library image. It has no
counterpart in the reference; it is checked against the existing entry point mtd_mapgen_batch_dev, which runs the same
kernel on cudaMalloc'ed buffers. Integer outputs: the tolerance is bit-exact, as everywhere in tests/.

Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

KAT_SEED = (0x1234567890ABCDEF, 0xFEDCBA0987654321)
W = H = 101
T = W * H
NWP = 4


def _generate(ctx, torch, dev, first, nmaps, cap=512):
    """the existing entry point: maps generated on the device + their path lengths / distances / aggregates"""
    d = dict(
        elev=torch.empty((nmaps, T), dtype=torch.int16, device=dev),
        flags=torch.empty((nmaps, T), dtype=torch.int16, device=dev),
        wp_tiles=torch.empty((nmaps, NWP), dtype=torch.int32, device=dev),
        tiles=torch.empty((nmaps, NWP, T), dtype=torch.int32, device=dev),
        dist=torch.empty((nmaps, NWP, T), dtype=torch.int32, device=dev),
        fullpath=torch.empty((nmaps, NWP, cap), dtype=torch.int32, device=dev),
        fullpath_len=torch.empty((nmaps, NWP), dtype=torch.int32, device=dev),
        full_distance=torch.empty((nmaps, NWP), dtype=torch.int32, device=dev),
        agg=torch.empty((nmaps, 3), dtype=torch.int32, device=dev),
    )
    ctx.mapgen_batch_dev(KAT_SEED[0], KAT_SEED[1], first, nmaps, {k: v.data_ptr() for k, v in d.items()}, W, H, 50, 50, 45, NWP, cap)
    ctx.synchronize()
    return d


def _noalloc(ctx, torch, dev, d, nmaps):
    out = dict(
        fullpath_len=torch.full((nmaps, NWP), -1, dtype=torch.int32, device=dev),
        full_distance=torch.full((nmaps, NWP), -1, dtype=torch.int32, device=dev),
        agg=torch.full((nmaps, 3), -1, dtype=torch.int32, device=dev),
    )
    torch.cuda.synchronize()
    before, free_before = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
    ctx.waypoints_paths_noalloc(W, H, nmaps, NWP, d["elev"].data_ptr(), d["flags"].data_ptr(), d["wp_tiles"].data_ptr(), out["fullpath_len"].data_ptr(), out["full_distance"].data_ptr(), out["agg"].data_ptr())
    ctx.synchronize()
    # nothing was allocated on the device: neither through torch nor behind its back (the box runs nothing else)
    assert torch.cuda.memory_allocated(dev) == before and torch.cuda.mem_get_info(dev)[0] >= free_before
    return out


def test_noalloc_paths_match_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from mazetd_b200 import capi

    dev = torch.device("cuda", 0)
    ctx = capi.Context(0)
    try:
        # two calls in a row on different maps and different counts: the second runs on the array as the first left it
        for first, nmaps in ((300, 12), (7, 5), (300, 12)):
            d = _generate(ctx, torch, dev, first, nmaps)
            out = _noalloc(ctx, torch, dev, d, nmaps)
            for k in ("fullpath_len", "full_distance", "agg"):
                assert np.array_equal(out[k].cpu().numpy(), d[k].cpu().numpy()), (first, nmaps, k)
            assert int(d["fullpath_len"].min()) > 0  # the comparison is about real paths
        # beyond the fixed scratch: a clear error, nothing launched
        d = _generate(ctx, torch, dev, 0, 13)
        launches = ctx.launches
        with pytest.raises(capi.MtdError) as e:
            _noalloc(ctx, torch, dev, d, 13)
        assert "MTD_NOALLOC_SCRATCH_BYTES" in str(e.value) and ctx.launches == launches
        assert 256 + 12 * NWP * T * 8 <= capi.Context.NOALLOC_SCRATCH_BYTES < 256 + 13 * NWP * T * 8
    finally:
        ctx.close()
