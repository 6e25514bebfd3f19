"""torchrun worker for tests/test_shard_gpu.py: the column-slab sharded filter vs (a) the dense-literal oracle on
the full config-1 loop (appends, associations, predicts) and (b) the single-GPU FULL path on a populated map
(must be bit-identical: same inputs, same arithmetic, same order)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def main():
    import torch
    import torch.distributed as dist
    import ekf_slam_b200 as E
    from ekf_slam_b200.synth import make_world, synthetic_spd_state, observe_from_truth
    from oracle import dense_literal as O
    local = int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    rank, world = dist.get_rank(), dist.get_world_size()

    def live(slab, c_lo, storage):
        """Mask of live elements of a slab: everything (FULL) or rows <= global column (SYM)."""
        if storage == 'full':
            return np.ones(slab.shape, dtype=bool)
        i = np.arange(slab.shape[0])[:, None]
        j = c_lo + np.arange(slab.shape[1])[None, :]
        return i <= j

    if os.environ.get('SHARD_CASE') == '80k':
        # BASELINE config 4: 80 000 landmarks (n = 160 003, P 205 GB dense) in SYM column slabs; x and sampled rows
        # of P against the matrix-free oracle; 1e-9 relative
        from bench import make_inputs
        from oracle.lowrank_lazy import LazyLowRankFilter
        N, k, T = int(os.environ.get('SHARD_LANDMARKS', '80000')), 8, int(os.environ.get('SHARD_STEPS', '2'))
        x, A, d, s, lms, us, obs = make_inputs(N, k, T)
        g = E.shard.create_sharded(E.EKF_SLAM_UC, dist, local, storage='sym', capacity_landmarks=N, defer_m=int(os.environ.get('SHARD_DEFER', '8')))
        g.set_state_lowrank(x, s, A, d)
        if os.environ.get('EKF_SHADOW_SLABS') == '1':
            assert g.flush_info()[0], 'continuous flush not active on this slab'
        o = LazyLowRankFilter(x, A, d, s, Rc=g.Rc)
        for t in range(T):
            g.predict(us[t]); o.predict(us[t])
            rows = obs[t]
            g.update_batch(rows[:, :2], np.column_stack([rows[:, 0] * g.Rc[0], rows[:, 1] * g.Rc[1]]), rows[:, 2].astype(np.int32))
            acts = o.measure_rows(rows)
            assert [i for _, i in acts] == [int(r[2]) for r in rows]
        _, _, c_lo, c_hi = g.shard_info()
        n = 3 + 2 * N
        c_hi = min(c_hi, n)
        ex = rel(g.x, o.x)
        rng = np.random.default_rng(5 + rank)
        sample = sorted(set([0, 1, 2, 3 + 2 * (int(obs[0][0][2]) - 1), c_lo] + list(rng.integers(0, c_hi, 12))))
        worst = 0.0
        for i in sample:
            j0 = max(c_lo, i)                                   # SYM slab: rows <= column are live
            if j0 >= c_hi:
                continue
            got = g.P_block(i, j0, 1, c_hi - j0)
            want = o.block([i], j0, c_hi)
            worst = max(worst, rel(got, want))
        print('rank %d 80k-case N=%d slab cols [%d,%d): rel err x %.2e, %d sampled rows of P worst %.2e' %
              (rank, N, c_lo, c_hi, ex, len(sample), worst), flush=True)
        assert ex <= 1e-9 and worst <= 1e-9, (rank, ex, worst)
        g.close()
        dist.barrier()
        dist.destroy_process_group()
        if rank == 0:
            print('SHARD_WORKER_OK', flush=True)
        return

    for storage in ('full', 'sym'):
        # (a) full loop, N = 20, 400 steps, both variants, deferred batches
        for cls, ocls, mode in ((E.EKF_SLAM, O.EKF_SLAM, 'prefix'), (E.EKF_SLAM_UC, O.EKF_SLAM_UC, 'nearest')):
            w = make_world(N=20, T=400, k=5, mode=mode)
            g = E.shard.create_sharded(cls, dist, local, storage=storage, capacity_landmarks=20, defer_m=8)
            o = ocls()
            for t in range(w.T):
                g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t))
                o.predict(w.u[t]); o.measure(None, w.u[t], w.stub(t))
            assert g.chosen == o.chosen
            slab, c_lo = g.P_slab()
            mk = live(slab, c_lo, storage)
            ex, eP = rel(g.x, o.x), rel(slab * mk, o.P[:, c_lo:c_lo + slab.shape[1]] * mk)
            assert ex <= 1e-9 and eP <= 1e-9, (cls.__name__, rank, ex, eP)
            assert np.array_equal(g.s, o.s)
            print('rank %d %s/%s loop: N=%d slab cols [%d,%d) rel err x %.2e P %.2e' %
                  (rank, cls.__name__, storage, g.N, c_lo, c_lo + slab.shape[1], ex, eP), flush=True)
            g.close()

        # (b) populated 1k map: sharded deferred batches vs single-GPU FULL immediate path, bitwise
        N = 1000
        x, A, d, s, lms = synthetic_spd_state(N, seed=3)
        g = E.shard.create_sharded(E.EKF_SLAM, dist, local, storage=storage, capacity_landmarks=N, defer_m=8)
        ref = E.EKF_SLAM(capacity_landmarks=N, storage=storage, device=local)
        for f in (g, ref):
            f.set_state_lowrank(x, s, A, d)
        rng = np.random.default_rng(11)
        for step in range(4):
            rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:11]
            Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
            for f in (g, ref):
                f.predict([0.3, 1.0])
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
        slab, c_lo = g.P_slab()
        Pref = ref.P
        assert np.array_equal(g.x, ref.x), rel(g.x, ref.x)
        mk = live(slab, c_lo, storage)
        assert np.array_equal(slab * mk, Pref[:, c_lo:c_lo + slab.shape[1]] * mk), rel(slab * mk, Pref[:, c_lo:c_lo + slab.shape[1]] * mk)
        print('rank %d populated 1k/%s: slab bit-identical to the single-GPU path' % (rank, storage), flush=True)
        g.close(); ref.close()

    # (c) cross-step deferral on SYM slabs (defer_m = 16: terms stay pending across predict, the exchanged P0 columns are
    # corrected by the pending terms on arrival): bitwise equal to the single-GPU defer-16 path, x read mid-run included
    N = 1000
    x, A, d, s, lms = synthetic_spd_state(N, seed=3)
    os.environ['EKF_LAZY_FLUSH'] = '1'       # same flush policy on both sides (default: lazy on slabs, eager on one GPU)
    g = E.shard.create_sharded(E.EKF_SLAM, dist, local, storage='sym', capacity_landmarks=N, defer_m=16)
    ref = E.EKF_SLAM(capacity_landmarks=N, storage='sym', device=local, defer_m=16)
    del os.environ['EKF_LAZY_FLUSH']
    assert g.pend_cap() == 16 and ref.pend_cap() == 16
    for f in (g, ref):
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(12)
    for step in range(9):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:(8, 11, 5, 8, 1, 8, 8, 3, 8)[step]]
        Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
        for f in (g, ref):
            f.predict([0.3, 1.0])
            if step == 4:
                f.update(rows[0, :2], Rd[0], int(rows[0, 2]))
            else:
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
        if step % 3 == 1:
            assert np.array_equal(g.x, ref.x), (step, rel(g.x, ref.x))
    slab, c_lo = g.P_slab()
    Pref = ref.P
    assert np.array_equal(g.x, ref.x), rel(g.x, ref.x)
    mk = live(slab, c_lo, 'sym')
    assert np.array_equal(slab * mk, Pref[:, c_lo:c_lo + slab.shape[1]] * mk), rel(slab * mk, Pref[:, c_lo:c_lo + slab.shape[1]] * mk)
    print('rank %d populated 1k/sym defer 16 (cross-step deferral on slabs): slab bit-identical to the single-GPU path' % rank, flush=True)
    g.close(); ref.close()
    # and the full loop with appends and associations against the oracle
    w = make_world(N=20, T=300, k=5, mode='nearest')
    g = E.shard.create_sharded(E.EKF_SLAM_UC, dist, local, storage='sym', capacity_landmarks=20, defer_m=16)
    o = O.EKF_SLAM_UC()
    for t in range(w.T):
        g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t))
        o.predict(w.u[t]); o.measure(None, w.u[t], w.stub(t))
    assert g.chosen == o.chosen
    slab, c_lo = g.P_slab()
    mk = live(slab, c_lo, 'sym')
    ex, eP = rel(g.x, o.x), rel(slab * mk, o.P[:, c_lo:c_lo + slab.shape[1]] * mk)
    assert ex <= 1e-9 and eP <= 1e-9, (rank, ex, eP)
    print('rank %d UC loop sym defer 16: rel err x %.2e P %.2e' % (rank, ex, eP), flush=True)
    g.close()
    # (c2) continuous flush on SYM slabs (the K3 pass of every slab out of place on its own stream, robot strip, window of
    # pending terms; forced at this size): same estimator as the single-GPU continuous flush -- x at checkpoints and the slab
    # to 1e-12 -- with ragged batches, a single update, x / slab reads mid-run (which settle), and the UC loop with appends
    # and associations against the oracle
    os.environ['EKF_SHADOW_MIN_N'] = '0'
    os.environ['EKF_SHADOW_SLABS'] = '1'          # off by default on slabs (slower than the lazy flush at 10k landmarks)
    N = 1000
    x, A, d, s, lms = synthetic_spd_state(N, seed=3)
    g = E.shard.create_sharded(E.EKF_SLAM, dist, local, storage='sym', capacity_landmarks=N, defer_m=16)
    ref = E.EKF_SLAM(capacity_landmarks=N, storage='sym', device=local, defer_m=16)
    for f in (g, ref):
        f.set_state_lowrank(x, s, A, d)
    rng = np.random.default_rng(12)
    sizes = (8, 11, 5, 8, 1, 8, 8, 3, 8, 16, 8, 8)
    for step in range(36):
        rows = observe_from_truth(lms, x[:3], rng.choice(N, 24, replace=False), rng)[:sizes[step % len(sizes)]]
        Rd = np.column_stack([rows[:, 0] * 0.01, rows[:, 1] * 5.0])
        for f in (g, ref):
            f.predict([0.3, 1.0])
            if step % 12 == 4:
                f.update(rows[0, :2], Rd[0], int(rows[0, 2]))
            else:
                f.update_batch(rows[:, :2], Rd, rows[:, 2].astype(np.int32))
        if step % 3 == 1:
            assert rel(g.x, ref.x) <= 1e-12, (step, rel(g.x, ref.x))
        if step == 17:
            slab, c_lo = g.P_slab()
            mk = live(slab, c_lo, 'sym')
            e17 = rel(slab * mk, ref.P[:, c_lo:c_lo + slab.shape[1]] * mk)
            assert e17 <= 1e-12, (rank, e17)
    slab, c_lo = g.P_slab()
    mk = live(slab, c_lo, 'sym')
    eP = rel(slab * mk, ref.P[:, c_lo:c_lo + slab.shape[1]] * mk)
    assert rel(g.x, ref.x) <= 1e-12 and eP <= 1e-12, (rank, rel(g.x, ref.x), eP)
    print('rank %d continuous flush on slabs vs single GPU: rel err x %.2e slab %.2e' % (rank, rel(g.x, ref.x), eP), flush=True)
    g.close(); ref.close()
    w = make_world(N=20, T=300, k=5, mode='nearest')
    g = E.shard.create_sharded(E.EKF_SLAM_UC, dist, local, storage='sym', capacity_landmarks=20, defer_m=16)
    o = O.EKF_SLAM_UC()
    for t in range(w.T):
        g.predict(w.u[t]); g.measure(None, w.u[t], w.stub(t))
        o.predict(w.u[t]); o.measure(None, w.u[t], w.stub(t))
    assert g.chosen == o.chosen
    slab, c_lo = g.P_slab()
    mk = live(slab, c_lo, 'sym')
    ex, eP = rel(g.x, o.x), rel(slab * mk, o.P[:, c_lo:c_lo + slab.shape[1]] * mk)
    assert ex <= 1e-9 and eP <= 1e-9, (rank, ex, eP)
    print('rank %d UC loop, continuous flush on slabs: rel err x %.2e P %.2e' % (rank, ex, eP), flush=True)
    g.close()
    del os.environ['EKF_SHADOW_MIN_N']
    del os.environ['EKF_SHADOW_SLABS']
    # (d) association mode B (Correspondence.m:74: Mahalanobis + signature) on slabs: every rank scores its own landmarks,
    # 16-byte candidates cross NVLink; indices identical to the single-GPU kernel and to oracle/structured.c, incl. a
    # duplicated landmark whose twin lives on another rank (tie -> first index) and a one-ulp twin
    from oracle import structured as S
    from ekf_slam_b200.synth import lowrank_to_dense
    for storage in ('sym', 'full'):
        N = 600
        x, A, d, s, lms = synthetic_spd_state(N, seed=5)
        P = lowrank_to_dense(A, d)
        for a, b, ulp in ((7, 580, False), (40, 300, True), (590, 12, False)):
            ia, ib = 3 + 2 * a, 3 + 2 * b
            x[ib:ib + 2] = x[ia:ia + 2]; lms[b] = lms[a]
            P[:3, ib:ib + 2] = P[:3, ia:ia + 2]; P[ib:ib + 2, :3] = P[ia:ia + 2, :3]
            P[ib:ib + 2, ib:ib + 2] = P[ia:ia + 2, ia:ia + 2]
            if ulp:
                x[ib] = np.nextafter(x[ib], np.inf)
        ones = np.ones(N)
        g = E.shard.create_sharded(E.EKF_SLAM_UC, dist, local, storage=storage, capacity_landmarks=N, assoc_mode='position')
        ref = E.EKF_SLAM_UC(capacity_landmarks=N, storage=storage, device=local, assoc_mode='position')
        for f in (g, ref):
            f.correspondence.s_thresh = 9.21
            f._sync_corr()
            f.set_state(x, P, ones)
        o = S.EKF_SLAM_UC(capacity_landmarks=N, assoc_mode=1)
        o.s_thresh = 9.21
        o.set_state(x, P, ones)
        rng = np.random.default_rng(2)
        ids = np.concatenate([[7, 580, 40, 300, 590, 12], rng.choice(N, 120, replace=False)])
        rows = observe_from_truth(lms, x[:3], ids, rng, seam_deg=0.0)
        n_new = 0
        for r in rows:
            z = np.array([r[0], r[1], 1.0])
            Rd = np.array([z[0] * 0.1, z[1] * 5.0])
            got = g.correspondence.estimateCorrespondence(z, np.diag(Rd))
            assert got == ref.correspondence.estimateCorrespondence(z, np.diag(Rd)) == o.associate(z, Rd), (rank, z, got)
            n_new += int(got[0])
        z = np.array([500.0, 170.0, 1.0])                # far from everything: gated out on every rank -> new landmark N + 1
        far = g.correspondence.estimateCorrespondence(z, np.diag([1e-3, 1e-3]))
        assert far == ref.correspondence.estimateCorrespondence(z, np.diag([1e-3, 1e-3])) == (True, N + 1), far
        print('rank %d mode B on slabs/%s: %d queries identical to the single-GPU kernel and the C oracle (%d new)' %
              (rank, storage, len(rows), n_new), flush=True)
        g.close(); ref.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print('SHARD_WORKER_OK', flush=True)


if __name__ == '__main__':
    main()
