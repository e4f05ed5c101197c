"""Multi-process host logic on CPU (gloo, world_size 2): session sharding, the packed
sufficient-statistic all-reduce of `parallel.gibbs_sweep`, and replicated parameter draws.
The per-shard statistics come from the oracle here (no GPU in this test); on the GPU box the
same pack/all_reduce/unpack code runs on the kernels' outputs over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as td
import torch.multiprocessing as mp

import kpms_helpers as H
from oracle import kpms_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _local_stats(d, s, p, Gm, L, K, lo, hi):
    sl = slice(lo, hi)
    F = O.ar_sufficient_stats(nlags=L, num_states=K, mask=d['mask'][sl], x=s['x'][sl], z=s['z'][sl])
    S_y, msum = O.obs_variance_stats(Y=d['Y'][sl], mask=d['mask'][sl], Cd=p['Cd'], v=s['v'][sl], h=s['h'][sl],
                                     x=s['x'][sl], s=s['s'][sl], Gamma=Gm)
    counts = O.count_transitions(K, d['mask'][sl][:, -s['z'].shape[1]:], s['z'][sl])
    t = torch.as_tensor
    return t(F), t(S_y), t(np.array([msum])), t(counts.astype(np.int64))


def _worker(rank, world, port, out):
    from keypointmoseq_b200 import gibbs as G, parallel
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        N, T, k, D, L, K = 5, 60, 6, 10, 3, 4
        d, s, p, Gm = H.problem(seed=9, N=N, T=T, k=k, D=D, L=L, K=K, ragged=True)
        hp = H.hyp(D, L, K)
        lo, hi = parallel.shard_range(N, rank, world)
        F, S_y, msum, counts = _local_stats(d, s, p, Gm, L, K, lo, hi)
        buf = parallel.pack_stats(F, S_y, msum, counts)
        td.all_reduce(buf, op=td.ReduceOp.SUM)
        F2, S2, m2, c2 = parallel.unpack_stats(buf, K, F.shape[1], k)
        # the two-stage reduction of the host-driven sweep must give the same sums
        F3, _, _, c3 = parallel.all_reduce_stats(F, None, None, counts)
        S3, m3 = parallel.all_reduce_obs_stats(S_y, msum)
        F4, S4, m4, c4 = parallel.all_reduce_stats(F, S_y, msum, counts)
        two_stage_ok = bool(torch.equal(F3, F2) and torch.equal(c3, c2) and torch.equal(S3, S2) and
                            torch.equal(m3.reshape(-1), m2.reshape(-1)) and torch.equal(F4, F2) and
                            torch.equal(S4, S2) and torch.equal(c4, c2))
        # verdict flags of the chunked HMM passes ride in the same all-reduce; shard offsets from sizes
        flags = torch.tensor([float(rank == 1), 1.0 - float(rank == 1), 1.0], dtype=torch.float64)
        F5, S5, m5, c5, fsum = parallel.all_reduce_stats(F, S_y, msum, counts, None, flags)
        flags_ok = bool(torch.equal(F5, F2) and torch.equal(c5, c2) and fsum.tolist() == [1.0, 1.0, 2.0])
        off = parallel.session_offset_of(hi - lo)
        off_ok = (off == lo)
        Ab, Q = G.sample_ar_params((3, 1), F2, nlags=L, nu_0=hp['nu_0'], S_0=torch.as_tensor(hp['S_0']),
                                   M_0=torch.as_tensor(hp['M_0']), K_0=torch.as_tensor(hp['K_0']))
        sig = G.sample_obs_variance((3, 1), S2, m2, nu_sigma=hp['nu_sigma'], sigmasq_0=hp['sigmasq_0'])
        gathered = [torch.zeros_like(Ab) for _ in range(world)]
        td.all_gather(gathered, Ab)
        if rank == 0:
            Ff, Sf, mf, cf = _local_stats(d, s, p, Gm, L, K, 0, N)
            out.put(dict(F_err=float((F2 - Ff).abs().max() / Ff.abs().max()), S_err=float((S2 - Sf).abs().max()),
                         m_ok=bool(torch.equal(m2, mf)), c_ok=bool(torch.equal(c2, cf)), two_stage_ok=two_stage_ok,
                         flags_ok=flags_ok, off_ok=off_ok,
                         same_params=bool(all(torch.equal(g, Ab) for g in gathered)),
                         finite=bool(torch.isfinite(Ab).all() and torch.isfinite(Q).all() and (sig > 0).all())))
    finally:
        td.destroy_process_group()


def test_shard_range_partitions_sessions():
    from keypointmoseq_b200 import parallel
    for n, w in [(1000, 8), (125, 8), (5, 2), (3, 4)]:
        spans = [parallel.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1


def test_packed_allreduce_and_replicated_draws_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res["F_err"] < 1e-13 and res["S_err"] < 1e-9
    assert res["m_ok"] and res["c_ok"] and res["same_params"] and res["finite"] and res["two_stage_ok"]
    assert res["flags_ok"] and res["off_ok"]


def test_shared_hmm_warmup_moves_identically_from_summed_verdicts():
    """The warm-up multiplier of the chunked HMM passes is a function of all-reduced verdict counts
    only, so every rank changes it at the same sweep."""
    from keypointmoseq_b200 import gibbs as G, parallel
    saved = dict(G._HMM_ADAPT)
    fresh = dict(boost=1, streak=0, need=8, probing=False)
    try:
        G._HMM_ADAPT.clear()
        G._HMM_ADAPT.update(fresh)
        parallel._apply_verdicts(1.0, 6.0, 8.0)          # one of eight ranks failed its check: a lone failure,
        assert G._HMM_ADAPT["boost"] == 1                # redone by that rank, the multiplier stays
        parallel._apply_verdicts(0.0, 8.0, 8.0)
        parallel._apply_verdicts(2.0, 6.0, 8.0)          # a second failure two sweeps later: systematic
        assert G._HMM_ADAPT["boost"] == 2 and G._HMM_ADAPT["streak"] == 0
        for _ in range(7):
            parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert G._HMM_ADAPT["boost"] == 2 and G._HMM_ADAPT["streak"] == 7
        parallel._apply_verdicts(0.0, 7.0, 8.0)          # one rank not comfortable: the streak starts over
        assert G._HMM_ADAPT["streak"] == 0
        for _ in range(8):
            parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert G._HMM_ADAPT["boost"] == 1 and G._HMM_ADAPT["probing"]
        before = dict(G._HMM_ADAPT)
        parallel._apply_verdicts(0.0, 0.0, 0.0)          # nothing reported (sequential passes): no change
        assert G._HMM_ADAPT == before
        parallel._apply_verdicts(1.0, 0.0, 8.0)          # the shorter warm-up fails: probe less often from now on
        assert (G._HMM_ADAPT["boost"], G._HMM_ADAPT["need"], G._HMM_ADAPT["probing"]) == (2, 32, False)
        for _ in range(31):
            parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert G._HMM_ADAPT["boost"] == 2
        parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert G._HMM_ADAPT["boost"] == 1 and G._HMM_ADAPT["probing"]
        for _ in range(8):                               # the probe survives: probation over, `need` kept
            parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert not G._HMM_ADAPT["probing"] and G._HMM_ADAPT["need"] == 32
        for _ in range(10):
            parallel._apply_verdicts(3.0, 0.0, 8.0)
        assert G._HMM_ADAPT["boost"] == G.HMM_MAX_BOOST
        G._HMM_ADAPT.update(frozen=True)                 # while a failed pass is being redone nothing moves
        parallel._apply_verdicts(0.0, 8.0, 8.0)
        assert G._HMM_ADAPT["streak"] == 0
    finally:
        G._HMM_ADAPT.clear()
        G._HMM_ADAPT.update(saved)


def test_kalman_warmup_multiplier_probes_below_one_and_backs_off():
    from keypointmoseq_b200 import gibbs as G
    saved = dict(G._LAT_ADAPT)
    tol = G.LATENTS_CHUNKING["tol"]
    try:
        G._LAT_ADAPT.update(boost=1, streak=0, need=3, probing=False)
        for _ in range(3):
            assert G.note_latents_check(1e-15)
        assert G._LAT_ADAPT["boost"] == 0.5 and G._LAT_ADAPT["probing"]      # comfortable three times: try half
        assert G.note_latents_check(0.02 * tol)                              # the shorter warm-up holds
        for _ in range(20):
            G.note_latents_check(1e-15)
        assert G._LAT_ADAPT["boost"] == 0.5                                  # never below the floor
        G._LAT_ADAPT.update(boost=0.5, streak=0, need=3, probing=True)
        assert not G.note_latents_check(10 * tol)                            # the probe fails: back to 1, re-probe later
        assert G._LAT_ADAPT["boost"] == 1 and G._LAT_ADAPT["need"] == 12
        assert not G.note_latents_check(float("nan"))                        # NaN is a failure, the second in a row
        assert G._LAT_ADAPT["boost"] == 2
    finally:
        G._LAT_ADAPT.clear()
        G._LAT_ADAPT.update(saved)
