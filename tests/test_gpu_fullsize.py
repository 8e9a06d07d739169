"""Parity at BASELINE.json's full sizes through size-independent properties (the oracle cannot finish
these shapes in seconds): planted-query recovery, tensor-core engine == exact engine on a subsample,
oracle check on a small subsample against the FULL dataset, KModes optimality of the returned labels."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    import tiler_b200 as tb
    from bench import synth_queries, synth_tileset  # the reference LCG (checked against oracle/gen.c on the CPU)
    tb.set_device(0)
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(1234)
    return dict(torch=torch, tb=tb, dev=dev, gen=gen, synth_tileset=synth_tileset, synth_queries=synth_queries)


def _render(torch, tiles, pal_i32, tid, hm, vm):
    """exact render of tile tid in orientation (hm, vm) through the palette -> $00BBGGRR pixels"""
    px = tiles[tid].view(-1, 8, 8).long()
    px = torch.where(hm.view(-1, 1, 1), px.flip(2), px)
    px = torch.where(vm.view(-1, 1, 1), px.flip(1), px)
    return pal_i32[px.reshape(-1, 64)].contiguous()


@pytest.mark.parametrize("ntiles,nq", [(16384, 1 << 20), (65536, 1 << 19)])  # C2 / C3 tilesets (N = 65 536 / 262 144 rows)
def test_planted_queries_recovered_at_full_n(env, oracle, ntiles, nq):
    torch, tb, dev, gen = env["torch"], env["tb"], env["dev"], env["gen"]
    tiles, pal = env["synth_tileset"](torch, ntiles, 16, gen)
    tid = torch.randint(0, ntiles, (nq,), generator=gen, device=dev)
    hm = torch.randint(0, 2, (nq,), generator=gen, device=dev).bool()
    vm = torch.randint(0, 2, (nq,), generator=gen, device=dev).bool()
    rgb = _render(torch, tiles, pal, tid, hm, vm)
    with tb.KDTree.from_tileset(tiles, pal.view(1, -1)) as t:
        idx = torch.empty(nq, dtype=torch.int32, device=dev)
        err = torch.empty(nq, dtype=torch.float32, device=dev)
        tb.reset_stats()
        t.frametiling_search(rgb, out_idx=idx, out_err=err)
        torch.cuda.synchronize()
        st = tb.stats()
        # the query IS a dataset row: squared distance exactly 0, and the returned row renders to the same pixels
        # (duplicate / symmetric tiles may legitimately return another row with identical pixels)
        assert float(err.abs().max()) == 0.0
        want = (tid.long() * 2 + vm.long()) * 2 + hm.long()  # row = ((p*T + t)*2 + vmir)*2 + hmir, main.pas:3902-3910
        diff = (idx.long() != want).nonzero().flatten()
        if len(diff):
            got = idx[diff].long()
            r2 = _render(torch, tiles, pal, got >> 2, (got & 1).bool(), ((got >> 1) & 1).bool())
            assert torch.equal(r2, rgb[diff])
            assert torch.all(got < want[diff])  # lowest row index among exact duplicates
        assert st["fallback_queries"] <= nq // 100
        # noisy queries: tensor-core engine == exact SIMT engine on a subsample, oracle on a smaller one
        m = 8192
        noisy = env["synth_queries"](torch, m, tiles, pal, gen)
        i_tc = torch.empty(m, dtype=torch.int32, device=dev)
        e_tc = torch.empty(m, dtype=torch.float32, device=dev)
        t.frametiling_search(noisy, out_idx=i_tc, out_err=e_tc)
        tb.set_engine(tb.ENGINE_SIMT)
        try:
            i_ex = torch.empty_like(i_tc)
            e_ex = torch.empty_like(e_tc)
            t.frametiling_search(noisy, out_idx=i_ex, out_err=e_ex)
        finally:
            tb.set_engine(tb.ENGINE_AUTO)
        torch.cuda.synchronize()
        assert torch.equal(i_tc, i_ex) and torch.equal(e_tc.view(torch.int32), e_ex.view(torch.int32))
        o = oracle
        sub = noisy[:256].cpu().numpy()
        ds = o.psyv_tileset_f32(tiles.cpu().numpy(), pal.cpu().numpy())
        oi, od, nxt = o.knn_f64(ds, o.psyv_queries_f32(sub), 1)
        solid = (nxt - od[:, 0]) > 1e-5 * np.maximum(od[:, 0], 1e-12)
        assert np.array_equal(i_tc[:256].cpu().numpy()[solid], oi[solid, 0])
        assert np.allclose(e_tc[:256].cpu().numpy(), od[:, 0], rtol=1e-5, atol=1e-7)


def test_kmodes_c4_properties(env, oracle):
    """1 048 576 points x 16 384 centroids: planted points get dissimilarity 0, every returned (label, dis) pair is
    consistent with a recomputation, and no sampled centroid beats the returned one."""
    torch, tb, dev, gen = env["torch"], env["tb"], env["dev"], env["gen"]
    npnt, k = 1 << 20, 1 << 14
    t2, _ = env["synth_tileset"](torch, npnt, 16, gen)
    flags = torch.stack([(t2 == z).sum(1) > 1 for z in range(16)], 1).to(torch.uint8)
    X = torch.cat([t2, flags], 1).contiguous()
    cent = X[:: npnt // k][:k].contiguous()
    clust = torch.empty(npnt, dtype=torch.int32, device=dev)
    dis = torch.empty(npnt, dtype=torch.int64, device=dev)
    tb.kmodes_assign(X, cent, tb.KM_ASM, out_clust=clust, out_dis=dis)
    torch.cuda.synchronize()
    planted = torch.arange(0, npnt, npnt // k, device=dev)[:k]
    assert torch.all(dis[planted] == 0)
    # a planted point's label is the LAST centroid equal to it (kmodes.pas:414-415)
    cl = clust[planted].long()
    assert torch.equal(cent[cl], X[planted]) and torch.all(cl >= torch.arange(k, device=dev))
    o = oracle
    sel = torch.randint(0, npnt, (4096,), generator=gen, device=dev)
    Xs, cs, ds_ = X[sel].cpu().numpy(), clust[sel].cpu().numpy(), dis[sel].cpu().numpy()
    cn = cent.cpu().numpy()
    for j in range(0, 4096, 64):  # recomputed dissimilarity of the returned pair
        assert o.dissim(cn[cs[j]], Xs[j], o.ASM) == ds_[j]
    wc, wd = o.kmodes_assign(Xs[:512], cn, o.ASM)  # full arg-min on a subsample against ALL centroids
    assert np.array_equal(wc, cs[:512]) and np.array_equal(wd.astype(np.int64), ds_[:512])


@pytest.mark.parametrize("nq", [(1 << 16) - 1, 1 << 16, (1 << 16) + 77, 200_003])
def test_coherence_order_is_invisible(env, oracle, nq):
    """Batches of >= 65 536 queries are searched in a colour-sorted order inside the library (qorder.cu) and
    scattered back: device-pointer and host-buffer calls must return, query for query, what the exact SIMT engine
    returns in the caller's order (ragged sizes on both sides of the threshold), and the oracle's answer on a
    subsample."""
    torch, tb, dev, gen = env["torch"], env["tb"], env["dev"], env["gen"]
    tiles, pal = env["synth_tileset"](torch, 2048, 16, gen)
    rgb = env["synth_queries"](torch, nq, tiles, pal, gen)
    with tb.KDTree.from_tileset(tiles, pal.view(1, -1)) as t:
        i_tc = torch.empty(nq, dtype=torch.int32, device=dev)
        e_tc = torch.empty(nq, dtype=torch.float32, device=dev)
        tb.set_engine(tb.ENGINE_TC)
        t.frametiling_search(rgb, out_idx=i_tc, out_err=e_tc)
        h_i, h_e = t.frametiling_search(rgb.cpu().numpy())          # host buffers: 2^19-query chunks
        tb.set_engine(tb.ENGINE_SIMT)
        i_ex = torch.empty(nq, dtype=torch.int32, device=dev)
        e_ex = torch.empty(nq, dtype=torch.float32, device=dev)
        t.frametiling_search(rgb, out_idx=i_ex, out_err=e_ex)
        tb.set_engine(tb.ENGINE_AUTO)
        torch.cuda.synchronize()
        assert torch.equal(i_tc, i_ex)
        assert torch.equal(e_tc.view(torch.int32), e_ex.view(torch.int32))
        assert np.array_equal(h_i, i_ex.cpu().numpy())
        assert np.array_equal(h_e.view(np.uint32), e_ex.cpu().numpy().view(np.uint32))
        # oracle on a strided subsample (positions spread over the whole batch)
        sel = torch.arange(0, nq, max(1, nq // 256), device=dev)
        ds = t.rows()
        q = oracle.psyv_queries_f32(rgb[sel].cpu().numpy())
        oi, od, nxt = oracle.knn_f64(ds, q, 1)
        solid = (nxt - od[:, 0]) > 1e-5 * np.maximum(od[:, 0], 1e-12)
        got_i, got_e = i_tc[sel].cpu().numpy(), e_tc[sel].cpu().numpy()
        assert np.array_equal(got_i[solid], oi[solid, 0])
        assert np.allclose(got_e, od[:, 0], rtol=1e-5, atol=1e-7)


def test_host_batch_larger_than_one_chunk_is_pipelined_and_consistent(env):
    """ann_kdtree_search_batch with host buffers cuts a batch into chunks of 2^20 queries and uploads chunk i + 1 on
    the copy stream under the search of chunk i (api.cu: search_batch_one).  Results of a ragged 2.2 M-query batch must
    equal separate calls on slices that straddle the chunk boundaries, and a brute-force check on a sample."""
    tb = env["tb"]
    rng = np.random.default_rng(77)
    n, dd, nq = 2048, 16, (1 << 21) + 100003
    ds = rng.standard_normal((n, dd)).astype(np.float32)
    q = rng.standard_normal((nq, dd)).astype(np.float32)
    with tb.KDTree.from_rows(ds) as t:
        for k in (1, 3):
            idx, err = t.search_batch(q, k=k)
            for lo, hi in [((1 << 20) - 300, (1 << 20) + 300), ((1 << 21) - 7, nq), (0, 1000)]:
                i2, e2 = t.search_batch(q[lo:hi], k=k)
                assert np.array_equal(idx[lo:hi], i2) and np.array_equal(err[lo:hi].view(np.uint32), e2.view(np.uint32))
            sample = np.concatenate([np.arange(0, 64), np.arange((1 << 20) - 32, (1 << 20) + 32), np.arange(nq - 64, nq)])
            d = ((q[sample, None, :].astype(np.float64) - ds[None, :, :].astype(np.float64)) ** 2).sum(-1)
            want = np.argsort(d, axis=1, kind="stable")[:, :k]
            gap = np.sort(d, axis=1)
            clear = (gap[:, k] - gap[:, k - 1]) > 1e-5 * np.maximum(gap[:, k - 1], 1e-12)
            if k == 1:
                assert np.array_equal(idx.reshape(nq, -1)[sample][clear, 0], want[clear, 0])
            assert np.allclose(err.reshape(nq, -1)[sample], np.sort(d, axis=1)[:, :k], rtol=1e-5, atol=1e-6)
