"""Host-side logic of the multi-GPU paths, on CPU: slab partition arithmetic (ekf_slab_range) and the
world_size-2 gloo plumbing that hands the NCCL id and the per-rank slab ranges around (no GPU, no NCCL traffic)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('storage', ['full', 'sym'])
@pytest.mark.parametrize('world', [1, 2, 4, 8])
@pytest.mark.parametrize('cap', [20, 1000, 10000, 80000])
def test_slab_ranges_partition_the_columns(cap, world, storage, lib):
    from ekf_slam_b200.shard import slab_range, owner_of_column
    n = 3 + 2 * cap
    r = [slab_range(cap, world, g, storage) for g in range(world)]
    assert r[0][0] == 0 and r[-1][1] >= n                       # rank 0 holds the robot columns; last slab ends at/after n
    for g in range(1, world):
        assert r[g][0] == r[g - 1][1]                           # contiguous, disjoint
        assert (r[g][0] - 3) % 2 == 0                           # landmark-aligned: both columns of a landmark on one rank
    if storage == 'full':
        w = [hi - lo for lo, hi in r]
        assert max(w) - min(w) <= 2 + 3 + 16                    # equal column counts up to alignment
    elif cap >= 1000:
        area = [(hi * (hi + 1) - lo * (lo + 1)) / 2 for lo, hi in ((lo, min(hi, n)) for lo, hi in r)]
        assert max(area) / (sum(area) / world) < 1.02           # equal live (upper-triangle) area within 2 %
    for col in (0, 2, 3, n // 2, n - 1):
        g = owner_of_column(cap, world, col, storage)
        assert r[g][0] <= col < r[g][1]


def test_slab_range_rejects_bad_queries(lib):
    from ekf_slam_b200.shard import slab_range
    for args in [(10, 0, 0), (10, 2, 2), (10, 2, -1), (0, 1, 0)]:
        with pytest.raises(ValueError):
            slab_range(*args)


WORKER = r'''
import os, sys
sys.path.insert(0, %r)
import numpy as np
import torch.distributed as dist
from ekf_slam_b200 import shard
dist.init_process_group('gloo')
rank, world = dist.get_rank(), dist.get_world_size()
nid = shard.nccl_id_via_torch(dist)                     # rank 0 draws the id, gloo broadcasts it
assert len(nid) == 128
ids = [None] * world
dist.all_gather_object(ids, nid)
assert all(i == ids[0] for i in ids), 'ranks disagree on the NCCL id'
assert any(b != 0 for b in ids[0]), 'id is all zeros'
for storage in ('full', 'sym'):
    mine = shard.slab_range(10000, world, rank, storage)
    allr = [None] * world
    dist.all_gather_object(allr, mine)
    assert allr[0][0] == 0 and all(allr[g][0] == allr[g - 1][1] for g in range(1, world)) and allr[-1][1] >= 20003
    # Monte-Carlo regime: filter f lives on rank f %% world -- a partition of 4096 filters with no overlap
    mine_f = set(range(rank, 4096, world))
    allf = [None] * world
    dist.all_gather_object(allf, len(mine_f))
    assert sum(allf) == 4096
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print('GLOO_WORKER_OK')
''' % ROOT


def test_world_size_2_gloo_plumbing(tmp_path, lib):
    script = tmp_path / 'gloo_worker.py'
    script.write_text(WORKER)
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
           '--master-addr', '127.0.0.1', '--master-port', '29631', str(script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0 and 'GLOO_WORKER_OK' in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
