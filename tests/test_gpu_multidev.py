"""ONE process driving several GPUs through the C ABI (tm_set_devices): datasets are built on the first device and
broadcast over NCCL, clip-sized host batches are cut into one contiguous shard per GPU, frame-sized and single-query
calls go to the least busy GPU.  Results must be bit-identical to the one-GPU call.  Reference call pattern: worker
threads of one encoder process on one shared handle (main.pas:972 -> 3992-4047, FFI extern.pas:63-67).
Needs a box with >= 2 GPUs (gpurun --gpus 2); skipped elsewhere."""
import gc
import os
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs in one box")


@pytest.fixture(scope="module")
def tb():
    os.environ.setdefault("TM_SHARD_MIN", "4096")  # shard C1-sized batches too (read once, at first use)
    import tiler_b200 as t
    yield t
    gc.collect()
    t.set_devices([0])


def _c1(o):
    """BASELINE config 1 (320x240 x 64 frames, 4k tiles) on the reference LCG"""
    pal, s = o.gen_palette(o.BASE_SEED)
    tiles, s = o.gen_tileset(s, 4096)
    rgb, s = o.gen_queries(s, 64 * 1200, tiles, pal)
    return pal, tiles, rgb


@needs2
def test_replicas_hold_identical_rows_and_sharded_search_is_bit_identical(tb, oracle):
    o = oracle
    pal, tiles, rgb = _c1(o)
    gc.collect()
    tb.set_devices([0])
    with tb.KDTree.from_tileset(tiles, pal) as t1:
        assert t1.replicas == 1
        want = t1.frametiling_search(rgb)
        want8 = t1.search_batch(o.psyv_queries_f32(rgb[:20000]), 8)
    nd = min(_ngpu(), 8)
    tb.set_devices(list(range(nd)))
    assert tb.get_devices() == list(range(nd)) and tb.comm_init_ms() > 0
    with tb.KDTree.from_tileset(tiles, pal) as t:
        assert t.replicas == nd
        r0 = t.rows(0)
        for r in range(1, nd):
            assert np.array_equal(r0.view(np.uint32), t.rows(r).view(np.uint32))  # what NCCL delivered
        tb.reset_stats()
        gi, ge = t.frametiling_search(rgb)                      # 76 800 queries: sharded over the devices
        assert np.array_equal(gi, want[0]) and np.array_equal(ge.view(np.uint32), want[1].view(np.uint32))
        for n in (1, 255, 4096 * nd + 777, 50001):              # ragged: one device, uneven shards
            gi, ge = t.frametiling_search(rgb[:n])
            assert np.array_equal(gi, want[0][:n]) and np.array_equal(ge.view(np.uint32), want[1][:n].view(np.uint32))
        gi, ge = t.search_batch(o.psyv_queries_f32(rgb[:20000]), 8)
        assert np.array_equal(gi, want8[0]) and np.array_equal(ge.view(np.uint32), want8[1].view(np.uint32))
        # legacy single-query symbols on a replicated handle
        q = o.psyv_queries_f32(rgb[:8])
        for i in range(8):
            idx, err = t.search(q[i])
            assert idx == want[0][i] and np.float32(err) == want[1][i]


@needs2
def test_frame_calls_from_many_threads_spread_over_devices(tb, oracle):
    o = oracle
    pal, tiles, rgb = _c1(o)
    gc.collect()
    tb.set_devices([0, 1])
    with tb.KDTree.from_tileset(tiles, pal) as t:
        want = t.frametiling_search(rgb)
        errors = []

        def worker(w):
            try:
                for rep in range(6):
                    f = (w * 6 + rep) % 64
                    n = 1200 if rep % 2 == 0 else 700 + 13 * w   # mixed sizes
                    gi, ge = t.frametiling_search(rgb[f * 1200:f * 1200 + n])
                    assert np.array_equal(gi, want[0][f * 1200:f * 1200 + n])
                    assert np.array_equal(ge.view(np.uint32), want[1][f * 1200:f * 1200 + n].view(np.uint32))
            except BaseException as e:  # noqa: BLE001
                errors.append((w, repr(e)))
        th = [threading.Thread(target=worker, args=(w,)) for w in range(16)]
        [x.start() for x in th]
        [x.join() for x in th]
        assert not errors, errors


@needs2
def test_device_list_cannot_change_under_live_handles(tb, oracle):
    from tiler_b200._lib import TMError
    gc.collect()
    tb.set_devices([0, 1])
    ds = np.random.default_rng(0).standard_normal((600, 192)).astype(np.float32)
    t = tb.KDTree.from_rows(ds)
    with pytest.raises(TMError, match="alive"):
        tb.set_devices([0])
    t.destroy()
    tb.set_devices([1])            # a single, non-default device
    with tb.KDTree.from_rows(ds) as t2:
        i, e = t2.search_batch(ds[:100], 1)
        assert np.array_equal(i[:, 0], np.arange(100))
    tb.set_devices([0])


@needs2
def test_device_pointer_calls_use_the_replica_on_that_device(tb, oracle):
    import torch
    o = oracle
    pal, tiles, rgb = _c1(o)
    gc.collect()
    tb.set_devices([0, 1])
    with tb.KDTree.from_tileset(tiles, pal) as t:
        want = t.frametiling_search(rgb[:30000])
        for dev in (0, 1):
            d = torch.device("cuda", dev)
            d_rgb = torch.from_numpy(rgb[:30000]).to(d)
            di = torch.empty(30000, dtype=torch.int32, device=d)
            de = torch.empty(30000, dtype=torch.float32, device=d)
            t.frametiling_search(d_rgb, out_idx=di, out_err=de)
            torch.cuda.synchronize(d)
            assert np.array_equal(di.cpu().numpy(), want[0])
            assert np.array_equal(de.cpu().numpy().view(np.uint32), want[1].view(np.uint32))
