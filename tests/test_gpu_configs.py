"""GPU (-m gpu): parity at the BASELINE.json config shapes that round 1 only timed, and the
multi-GPU data plane.

  * configs[2] (500 x 100k) and configs[4] (1000 x 200k): >= 200 mixed NNI/SPR/TBR proposals made
    by the host operators (the reference's mutate_*, /root/reference/src/TreeOperations.c:160-541),
    scored with lvb_gpu_score_batch, a subset committed, re-roots in between — every length, and
    after every commit every per-node count and sampled state rows, against the oracle.
  * configs[2] trajectory prefix: the unmodified reference binary run here under a wall-clock
    limit, `lvb_gpu -x <same count>`: identical changeAccepted.tsv lines.
  * NCCL: the all-reduce calls inside commit / score_batch / nni_neighbourhood / tree_changes with a
    communicator attached — one rank always, two ranks on two GPUs when two are visible.
  * the bounded dependency wait (fault injection) and DNAToBinary on a second context.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import cli_harness as H
import oracles
from lvb_b200 import Engine, GpuError, Topology, host, sharding, synthetic

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def tiny_msa(n: int):
    """a host alignment with n taxa: the tree operators only need n"""
    rng = np.random.default_rng(7)
    rows = [bytes(b"ACGT"[i] for i in rng.integers(0, 4, size=8)) for _ in range(n)]
    return host.HostMSA.from_rows(rows)


@pytest.mark.parametrize("n,m,structured,keeping", [(500, 100_000, True, False), (1000, 200_000, False, False),
                                                    (500, 100_000, True, True), (1000, 200_000, False, True)])
def test_proposals_at_config_shapes(n, m, structured, keeping):
    """keeping: the accepted proposal of a round is scored AND kept (lvb_gpu_score_keep) and accepted with
    lvb_gpu_accept_kept instead of being committed — in two rounds of three; the third commits, so that commits
    run on trees whose rows sit in both banks"""
    nwords = (m + 15) // 16
    enc = synthetic.packed(n, m, seed=1, p=0.02) if structured else synthetic.random_packed(n, nwords, seed=1)
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=1)
        t = eng.tree_create()
        h = tiny_msa(n)
        h.rinit(4242)
        cur = h.tree().random()
        props = [h.tree() for _ in range(6)]
        _, left, right, _ = cur.get()
        want, rows, chg = oracles.oracle_getplen(enc, left, right, cur.root)
        assert eng.commit(t, Topology(left, right, cur.root)) == want
        assert np.array_equal(eng.changes(t)[n:], chg[n:])
        cur.mark_clean()
        rng = np.random.default_rng(11)
        kinds = ("nni", "spr", "tbr")
        scored = committed = 0
        longest = 0
        for rnd in range(40):
            if rnd and rnd % 13 == 0:
                # arbreroot (TreeOperations.c:639-655): only the old-root..new-root path changes
                newroot, path = cur.arbreroot()
                _, left, right, _ = cur.get()
                rows[path, 0] = 0
                want, rows, chg = oracles.oracle_getplen(enc, left, right, newroot, rows, chg)
                assert eng.commit(t, Topology(left, right, newroot, path)) == want, ("reroot", rnd)
                cur.mark_clean()
            batch, wants, keep = [], [], None
            pick = int(rng.integers(0, 6))
            for k in range(6):
                prop = props[k].mutate(kinds[(rnd + k) % 3], cur)
                _, l2, r2, dflags = prop.get()
                dirty = np.nonzero(dflags)[0]
                longest = max(longest, len(dirty))
                saved_rows, saved_chg = rows[dirty].copy(), chg[dirty].copy()
                rows[dirty, 0] = 0
                w, rows, chg = oracles.oracle_getplen(enc, l2, r2, prop.root, rows, chg)
                batch.append(Topology(l2, r2, prop.root, dirty))
                wants.append(w)
                if k == pick:
                    keep = (l2.copy(), r2.copy(), dirty, rows[dirty].copy(), chg[dirty].copy(), w)
                rows[dirty], chg[dirty] = saved_rows, saved_chg  # the proposal is not applied
            l2, r2, dirty, new_rows, new_chg, w = keep
            if keeping and rnd % 3 != 2:
                got, ticket = eng.score_keep([t] * len(batch), batch, [1 if k == pick else 0 for k in range(6)])
                assert got.tolist() == wants, rnd
                eng.accept_kept(ticket, [pick])
            else:
                got = eng.score_batch([t] * len(batch), batch)
                assert got.tolist() == wants, rnd
                # accept one of them: commit on the device, apply on the host tree and in the oracle's rows
                assert eng.commit(t, Topology(l2, r2, cur.root, dirty)) == w, rnd
            scored += len(batch)
            rows[dirty], chg[dirty] = new_rows, new_chg
            cur, props[pick] = props[pick], cur  # SwapTrees (TreeOperations.c:1062-1080)
            cur.mark_clean()
            committed += 1
            if rnd % 8 == 0:
                assert np.array_equal(eng.changes(t)[n:], chg[n:]), rnd
                for v in [int(dirty[0]), int(dirty[-1]), n + int(rng.integers(0, n - 3))]:
                    assert np.array_equal(eng.download_row(t, v), rows[v]), (rnd, v)
        assert np.array_equal(eng.changes(t)[n:], chg[n:])
        assert scored >= 200 and committed >= 40 and longest >= 20
    finally:
        eng.close()


@pytest.mark.skipif(not os.path.exists(oracles.REF_BIN), reason="oracle/_ref/lvb_ref not on this box")
def test_config3_trajectory_prefix_vs_reference_binary():
    """BASELINE.json configs[2] shape, one chain: the reference (`lvb -v -p 1`) for 45 s of wall
    clock, then lvb_gpu cut off just after the same number of rearrangements (-x stops before the line of its last one): the first K lines of
    changeAccepted.tsv (operator, accepted, length, temperature, stack size per proposal) agree."""
    with tempfile.TemporaryDirectory() as td:
        os.makedirs(os.path.join(td, "ref"))
        os.makedirs(os.path.join(td, "gpu"))
        synthetic.write_phylip(os.path.join(td, "ref", "infile"), synthetic.survey_alignment(500, 100_000, 1, 0.02))
        os.link(os.path.join(td, "ref", "infile"), os.path.join(td, "gpu", "infile"))
        try:
            subprocess.run([oracles.REF_BIN, "-s", "12345", "-a", "1", "-p", "1", "-v"], cwd=os.path.join(td, "ref"),
                           capture_output=True, timeout=45)
        except subprocess.TimeoutExpired:
            pass
        ref_lines = open(os.path.join(td, "ref", "changeAccepted.tsv"), "rb").read().split(b"\n")[:-1]  # complete lines
        k = len(ref_lines) - 1  # without the header
        assert k >= 500, k
        p = subprocess.run([H.CLI, "-s", "12345", "-a", "1", "-v", "-x", str(k + 2)], cwd=os.path.join(td, "gpu"),
                           capture_output=True, text=True, timeout=600)
        assert p.returncode == 0, p.stdout[-800:]
        gpu_lines = open(os.path.join(td, "gpu", "changeAccepted.tsv"), "rb").read().split(b"\n")
        assert gpu_lines[: k + 1] == ref_lines
        assert "kernel launches" in p.stdout


def _site_block_reference(enc, left, right, root, cands):
    want, _, chg = oracles.oracle_getplen(enc, left, right, root)
    lens = []
    for c in cands:
        l2, r2 = getattr(c, "full", (c.left, c.right))  # bench.nni_candidates: overlays on the current tree's arrays
        lens.append(oracles.oracle_getplen(enc, l2, r2, c.root)[0])
    return want, chg, lens


def _nni_candidates(left, right, root, n, count, seed):
    sys.path.insert(0, ROOT)
    import bench

    return bench.nni_candidates(left, right, root, n, count, seed)


@pytest.mark.parametrize("route", ["p2p", "nccl"])
def test_nccl_single_rank_communicator(monkeypatch, route):
    """the all-reduce of the site-block mode (commit, score_batch, nni_neighbourhood, tree_changes)
    executed with a one-rank communicator, through both routes — the peer-memory kernel (default) and
    ncclAllReduce (LVB_GPU_P2P=0): same numbers as without a communicator"""
    monkeypatch.setenv("LVB_GPU_P2P", "1" if route == "p2p" else "0")
    n, m = 150, 20_000
    enc = synthetic.packed(n, m, seed=5)
    left, right, root = synthetic.random_topology(n, seed=9)
    topo = Topology(left, right, root)
    cands, index = _nni_candidates(left, right, root, n, 24, seed=2)
    want, chg, lens = _site_block_reference(enc, left, right, root, cands)
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=2)
        eng.comm_init(1, 0, Engine.comm_unique_id())
        t, scratch = eng.tree_create(), eng.tree_create()
        assert eng.commit(t, topo) == want
        assert np.array_equal(eng.changes(t)[n:], chg[n:])
        assert eng.score_batch([t] * len(cands), cands).tolist() == lens
        nn = eng.nni_neighbourhood(t, scratch, topo)
        assert nn[np.asarray(index)].tolist() == lens
        # an accepted move, then the rest of the neighbourhood again
        assert eng.commit(t, cands[3]) == lens[3]
        assert eng.rerun_last_batch(1).tolist() == [lens[3]]
    finally:
        eng.close()


def _gpu_count() -> int:
    import ctypes

    from lvb_b200 import load_library

    k = ctypes.c_int(0)
    load_library().lvb_gpu_device_count(ctypes.byref(k))
    return k.value


@pytest.mark.skipif(_gpu_count() < 2, reason="needs two GPUs (the site-block path with two ranks)")
@pytest.mark.parametrize("route", ["p2p", "nccl"])
def test_nccl_two_ranks_site_blocks(route):
    """BASELINE.json configs[3] in small on two GPUs: each rank holds half of the words of every row,
    lengths are all-reduced inside the C-ABI (peer-memory kernel, or ncclAllReduce with LVB_GPU_P2P=0);
    both ranks must report the oracle's numbers"""
    env = dict(os.environ, LVB_GPU_P2P="1" if route == "p2p" else "0", LVB_GPU_PROFILE="1")
    with tempfile.TemporaryDirectory() as td:
        procs = [subprocess.Popen([sys.executable, os.path.join(HERE, "run_siteblock_rank.py"), str(r), "2", td],
                                  stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env) for r in range(2)]
        outs = [p.communicate(timeout=600) for p in procs]
        for p, (o, e) in zip(procs, outs):
            assert p.returncode == 0, e[-2000:]
        res = [json.loads(o.strip().splitlines()[-1]) for o, _ in outs]
        for _, e in outs:  # the route that was asked for is the one that ran
            assert ("through peer memory" in e) == (route == "p2p"), e[-600:]
    n, m = 300, 60_000
    enc = synthetic.packed(n, m, seed=5)
    left, right, root = synthetic.random_topology(n, seed=9)
    cands, index = _nni_candidates(left, right, root, n, 16, seed=2)
    want, chg, lens = _site_block_reference(enc, left, right, root, cands)
    for r in res:
        assert r["commit"] == want and r["score"] == lens and r["nni_at_index"] == lens
        assert r["changes"] == chg[n:].tolist()
    assert res[0]["words"] + res[1]["words"] == (m + 15) // 16


def test_bounded_dependency_wait(monkeypatch):
    """a piece that waits for rows nobody will store gives up after LVB_GPU_SPIN_LIMIT polls and the
    call fails loudly; the context recovers (fault injected through LVB_GPU_FAULT=dep)"""
    n, m = 300, 40_000
    enc = synthetic.packed(n, m, seed=3)
    left, right, root = synthetic.random_topology(n, seed=4)
    topo = Topology(left, right, root)
    want = oracles.oracle_getplen(enc, left, right, root)[0]
    monkeypatch.setenv("LVB_GPU_FAULT", "dep")
    monkeypatch.setenv("LVB_GPU_SPIN_LIMIT", "8192")
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=1)
        t = eng.tree_create()
        with pytest.raises(GpuError, match="gave up waiting"):
            eng.commit(t, topo)
        with pytest.raises(GpuError):  # the trees must be evaluated from scratch after a device error
            eng.score_batch([t], [Topology(left, right, root, [])])
        assert eng.commit(t, topo) == want  # the fault was injected once; the context works again
        assert eng.score_batch([t], [Topology(left, right, root, [])]).tolist() == [want]
    finally:
        eng.close()


def test_pack_on_every_device():
    """DNAToBinary on the device needs its symbol table in the __constant__ memory of EVERY GPU a
    process uses (it used to be installed on the first one only)"""
    rng = np.random.default_rng(3)
    rows = [bytes(b"ACGTUYRWSKMBDHVNX?-"[i] for i in rng.integers(0, 19, size=333)) for _ in range(12)]
    want = oracles.oracle_pack(rows)
    for dev in range(max(1, min(2, _gpu_count()))):
        for _ in range(2):  # two contexts per device
            eng = Engine(dev)
            try:
                eng.pack(rows, max_trees=1)
                got = np.stack([eng.download_row(0, v) for v in range(len(rows))])
                assert np.array_equal(got, want), dev
            finally:
                eng.close()


def test_commit_submit_and_wait():
    """lvb_gpu_commit_batch_submit + lvb_gpu_batch_wait: several whole-tree evaluations in flight
    (compile of tree k+1 under the kernel of tree k) give the lengths of one-at-a-time commits"""
    n, m = 200, 30_000
    enc = synthetic.packed(n, m, seed=8)
    topos = [Topology(*synthetic.random_topology(n, seed=20 + s)) for s in range(9)]
    want = [oracles.oracle_getplen(enc, tp.left, tp.right, tp.root)[0] for tp in topos]
    eng = Engine(0)
    try:
        eng.upload(enc, max_trees=2)
        t0, t1 = eng.tree_create(), eng.tree_create()
        got, inflight = [], []
        for s, tp in enumerate(topos):
            inflight.append(eng.commit_batch_submit([t0 if s % 2 == 0 else t1], [tp]))
            if len(inflight) > 3:
                got.append(int(eng.batch_wait(inflight.pop(0))[0]))
        while inflight:
            got.append(int(eng.batch_wait(inflight.pop(0))[0]))
        assert got == want
        # the trees hold the last topologies committed to them
        _, _, chg = oracles.oracle_getplen(enc, topos[8].left, topos[8].right, topos[8].root)
        assert np.array_equal(eng.changes(t0)[n:], chg[n:])
        assert eng.score_batch([t1], [Topology(topos[7].left, topos[7].right, topos[7].root, [])]).tolist() == [want[7]]
    finally:
        eng.close()
