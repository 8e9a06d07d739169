"""Multi-GPU FrameTiling: query-sharded, tileset replicated.

The reference fans DoFrameTiling out one frame per worker thread over a shared, read-only
dataset (main.pas:972 -> 3992-4047).  Here the workers are one process per GPU
(torch.distributed; NCCL on GPUs, gloo in the CPU tests): every rank owns a contiguous range
of whole frames, the packed tileset is broadcast once from rank 0, each rank searches its
range with the local library, and (idx, err) are gathered to rank 0.  There is no exchange
inside the search, so there is no collective on the data path besides those two.

torch is imported lazily: it is plumbing for the process group, not part of the product path.
"""
from . import ann


def frame_shards(n_frames, positions_per_frame, world):
    """Contiguous whole-frame ranges, sizes differing by at most one frame.

    Returns [(q_begin, q_end)] * world in query units (frame * positions_per_frame)."""
    if world < 1 or n_frames < 0 or positions_per_frame < 0:
        raise ValueError("frame_shards: bad arguments")
    base, extra = divmod(n_frames, world)
    out, f0 = [], 0
    for r in range(world):
        f1 = f0 + base + (1 if r < extra else 0)
        out.append((f0 * positions_per_frame, f1 * positions_per_frame))
        f0 = f1
    return out


def point_shards(npoints, world):
    """Contiguous point ranges for the KModes assignment pass (points sharded, centroids replicated)."""
    return frame_shards(npoints, 1, world)


class ShardedFrameTiling:
    """One instance per rank.  `search_fn(tree_args, rgb) -> (idx, err)` defaults to the CUDA library;
    the gloo tests inject the CPU oracle instead (tests may use the oracle, the product never does)."""

    def __init__(self, group=None, search_fn=None, build_fn=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._search = search_fn or self._cuda_search
        self._build = build_fn or self._cuda_build
        self.tree = None

    # -- default (product) implementation: libtmann through the C ABI -------------------------
    @staticmethod
    def _cuda_build(palpix, pals, tile_flags):
        return ann.KDTree.from_tileset(palpix, pals, tile_flags)

    @staticmethod
    def _cuda_search(tree, rgb, out_idx=None, out_err=None):
        return tree.frametiling_search(rgb, out_idx=out_idx, out_err=out_err)

    # -- collectives ---------------------------------------------------------------------------
    def broadcast_tileset(self, palpix, pals, tile_flags=None, src=0):
        """One-time broadcast of the tileset (rank `src`'s tensors win), then the local dataset build.
        Tensors must be preallocated with identical shapes on every rank (torch tensors, any device)."""
        if self.world > 1:
            self.dist.broadcast(palpix, src, group=self.group)
            self.dist.broadcast(pals, src, group=self.group)
            if tile_flags is not None:
                self.dist.broadcast(tile_flags, src, group=self.group)
        self.tree = self._build(palpix, pals, tile_flags)
        return self.tree

    def search_local(self, rgb_local, out_idx=None, out_err=None):
        return self._search(self.tree, rgb_local, out_idx, out_err)

    def gather_results(self, idx_local, err_local, sizes, dst=0):
        """Gather the per-rank (idx, err) shards to rank `dst`; `sizes` = queries per rank.
        Equal shards use dist.gather; ragged shards are padded to the largest one."""
        import torch
        if self.world == 1:
            return idx_local, err_local
        m = max(sizes)
        if idx_local.shape[0] != m:  # ragged: pad to the common length
            pi = torch.zeros(m, dtype=idx_local.dtype, device=idx_local.device)
            pe = torch.zeros(m, dtype=err_local.dtype, device=err_local.device)
            pi[: idx_local.shape[0]] = idx_local
            pe[: err_local.shape[0]] = err_local
            idx_local, err_local = pi, pe
        gi = ge = None
        if self.rank == dst:
            gi = [torch.empty_like(idx_local) for _ in range(self.world)]
            ge = [torch.empty_like(err_local) for _ in range(self.world)]
        self.dist.gather(idx_local, gi, dst, group=self.group)
        self.dist.gather(err_local, ge, dst, group=self.group)
        if self.rank != dst:
            return None, None
        return (torch.cat([g[:n] for g, n in zip(gi, sizes)]), torch.cat([g[:n] for g, n in zip(ge, sizes)]))

    def close(self):
        if self.tree is not None and hasattr(self.tree, "destroy"):
            self.tree.destroy()
        self.tree = None


class ShardedKModesAssign:
    """KModes assignment pass over several GPUs: points sharded, centroids replicated.

    The reference thread-pools DoGMMD over points with shared, read-only centroids (kmodes.pas:825-835,
    984-997).  One instance per rank: the centroids are broadcast from rank `src` (they change every iteration
    on the Pascal side), each rank assigns its contiguous point range with the local library and (clust, dis)
    are gathered to rank `dst`.  `assign_fn(X_local, centroids) -> (clust, dis)` defaults to the CUDA library;
    the gloo tests inject the CPU oracle."""

    def __init__(self, group=None, assign_fn=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self._assign = assign_fn or self._cuda_assign

    @staticmethod
    def _cuda_assign(X_local, centroids):
        import torch
        clust = torch.empty(X_local.shape[0], dtype=torch.int32, device=X_local.device)
        dis = torch.empty(X_local.shape[0], dtype=torch.int64, device=X_local.device)
        ann.kmodes_assign(X_local, centroids, ann.KM_ASM, out_clust=clust, out_dis=dis)
        return clust, dis

    def assign(self, X_local, centroids, sizes, src=0, dst=0):
        """X_local: this rank's rows [n_r][80] (torch uint8); centroids: [k][80], preallocated on every rank;
        sizes: points per rank (point_shards).  Returns (clust, dis) for all points on rank dst, else (None, None)."""
        import torch
        if self.world > 1:
            self.dist.broadcast(centroids, src, group=self.group)
        clust, dis = self._assign(X_local, centroids)
        if self.world == 1:
            return clust, dis
        m = max(sizes)
        pc = torch.full((m,), -1, dtype=clust.dtype, device=clust.device)
        pd = torch.zeros(m, dtype=dis.dtype, device=dis.device)
        pc[: clust.shape[0]] = clust
        pd[: dis.shape[0]] = dis
        gc = gd = None
        if self.rank == dst:
            gc = [torch.empty_like(pc) for _ in range(self.world)]
            gd = [torch.empty_like(pd) for _ in range(self.world)]
        self.dist.gather(pc, gc, dst, group=self.group)
        self.dist.gather(pd, gd, dst, group=self.group)
        if self.rank != dst:
            return None, None
        return (torch.cat([g[:n] for g, n in zip(gc, sizes)]), torch.cat([g[:n] for g, n in zip(gd, sizes)]))
