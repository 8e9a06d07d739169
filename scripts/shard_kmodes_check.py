#!/usr/bin/env python
"""torchrun check of ShardedKModesAssign over NCCL: the gathered assignment equals the one-GPU assignment."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import tiler_b200 as tb
    from tiler_b200.shard import ShardedKModesAssign, point_shards
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    tb.set_device(local)
    dist.init_process_group("nccl")
    dev = torch.device("cuda", local)
    g = torch.Generator(device=dev).manual_seed(5)
    X = torch.randint(0, 16, (200003, 80), dtype=torch.uint8, device=dev, generator=g)
    C = torch.randint(0, 16, (4096, 80), dtype=torch.uint8, device=dev, generator=g)
    if rank != 0:
        C.zero_()
    shards = point_shards(X.shape[0], world)
    a, b = shards[rank]
    ska = ShardedKModesAssign()
    gc, gd = ska.assign(X[a:b].contiguous(), C, [q - p for p, q in shards])
    if rank == 0:
        wc = torch.empty(X.shape[0], dtype=torch.int32, device=dev)
        wd = torch.empty(X.shape[0], dtype=torch.int64, device=dev)
        tb.kmodes_assign(X, C, tb.KM_ASM, out_clust=wc, out_dis=wd)
        torch.cuda.synchronize()
        assert torch.equal(gc, wc) and torch.equal(gd, wd)
        print("sharded kmodes ok:", world, "ranks,", X.shape[0], "points")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
