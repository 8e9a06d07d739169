"""bench.py generates its inputs with the reference's LCG (RandInt, kmodes.pas:82-86) evaluated with
torch ops; oracle/gen.c is the scalar restatement of the same generator.  They must agree draw for draw, so that a
future FPC build of the reference can regenerate the benchmark inputs from the seeds alone."""
import numpy as np
import torch

import bench


def test_scalar_lcg_matches_oracle(oracle):
    s = bench.BASE_SEED
    so = oracle.BASE_SEED
    for rng in (2, 16, 256, 1 << 24, 65536, 13):
        s, v = bench.lcg_scalar(s, rng)
        vo, so = oracle.randint(rng, so)
        assert (v, s) == (vo, so)


def test_palette_tileset_queries_match_oracle(oracle):
    o = oracle
    cpu = torch.device("cpu")
    seed = (bench.BASE_SEED + 1) & bench.M32
    pal, s1 = bench.lcg_palette(seed, 16)
    opal, os1 = o.gen_palette(seed)
    assert pal == [int(x) for x in opal] and s1 == os1
    tiles, s2 = bench.lcg_tileset(torch, s1, 777, 16, cpu)
    otiles, os2 = o.gen_tileset(os1, 777)
    assert s2 == os2
    assert np.array_equal(tiles.numpy(), otiles)
    rgb, _ = bench.lcg_queries(torch, s2, 0, 3000, tiles, pal, cpu)
    orgb, _ = o.gen_queries(os2, 3000, otiles, opal)
    assert np.array_equal(rgb.numpy(), orgb)
    # a shard of the global query range is the same slice (multi-GPU ranks generate their own shard)
    part, _ = bench.lcg_queries(torch, s2, 1000, 1700, tiles, pal, cpu)
    assert np.array_equal(part.numpy(), orgb[1000:1700])


def test_kmodes_rows_match_oracle(oracle):
    cpu = torch.device("cpu")
    tiles, _ = bench.lcg_tileset(torch, 12345, 500, 16, cpu)
    X = bench.kmodes_rows_torch(torch, tiles)
    assert np.array_equal(X.numpy(), oracle.kmodes_rows(tiles.numpy()))


def test_workloads_cover_baseline_configs():
    import json
    import os
    cfgs = json.load(open(os.path.join(bench.ROOT, "BASELINE.json")))["configs"]
    assert sorted(w["cfg"] for w in bench.WORKLOADS.values()) == list(range(len(cfgs)))
    assert bench.WORKLOADS["c3_1080p_1000f_64k"]["frames"] * bench.WORKLOADS["c3_1080p_1000f_64k"]["pos"] == 32_400_000
    assert bench.WORKLOADS["c5_4k_2000f_256k_k8"]["tiles"] * 4 == 1_048_576
