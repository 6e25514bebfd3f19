"""Host plumbing for the column-slab sharded filter: one process per GPU (torchrun), torch.distributed only
carries the 128-byte NCCL id; the data path (exchange of the <= 19 slot columns per batch) is a peer-memory scatter
over NVLink inside libekfslam_b200.so, NCCL collectives as the fallback (EKF_SHARD_EXCHANGE=nccl)."""
import ctypes as ct

import numpy as np

from . import _lib as L


def slab_range(capacity_landmarks, world, rank, storage='full'):
    """Columns [c_lo, c_hi) of P stored by `rank` (pure host arithmetic, no GPU needed).
    'full': equal column counts; 'sym': equal live (upper-triangle) area."""
    lo, hi = ct.c_int32(), ct.c_int32()
    st = L.STORAGE_SYM if storage in ('sym', L.STORAGE_SYM) else L.STORAGE_FULL
    rc = L.load().ekf_slab_range(int(capacity_landmarks), int(world), int(rank), st, ct.byref(lo), ct.byref(hi))
    if rc != L.EKF_OK:
        raise ValueError('bad slab query (capacity=%r world=%r rank=%r)' % (capacity_landmarks, world, rank))
    return lo.value, hi.value


def owner_of_column(capacity_landmarks, world, col, storage='full'):
    for g in range(world):
        lo, hi = slab_range(capacity_landmarks, world, g, storage)
        if lo <= col < hi:
            return g
    raise ValueError(col)


def nccl_id_via_torch(dist, device=None):
    """Rank 0 draws an NCCL unique id; torch.distributed (any backend) hands it to the other ranks."""
    import torch
    buf = (ct.c_uint8 * 128)()
    if dist.get_rank() == 0:
        L.check(None, L.load().ekf_nccl_unique_id(buf))
    t = torch.tensor(list(bytes(buf)), dtype=torch.uint8)
    if device is not None:
        t = t.to(device)
    dist.broadcast(t, src=0)
    return bytes(t.cpu().numpy().astype(np.uint8).tobytes())


def create_sharded(cls, dist, device_index, storage='full', **kw):
    """cls = EKF_SLAM / EKF_SLAM_UC.  Collective: every rank of `dist` calls it with the same arguments."""
    import torch
    dev = torch.device('cuda', device_index) if dist.get_backend() == 'nccl' else None
    nid = nccl_id_via_torch(dist, dev)
    return cls(device=device_index, rank=dist.get_rank(), world=dist.get_world_size(), nccl_id=nid,
               storage=storage, **kw)
