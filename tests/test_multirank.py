"""The N > 1 paths on CPU: two `gloo` ranks, each with the oracle-backed mock device
(tests/mock_gpu.c) behind the C-ABI.

  * site blocks (BASELINE.json configs[3]): lvb_b200.sharding.word_block cuts the words of every
    row over the ranks, each rank evaluates the same trees on its block, the partial lengths are
    all-reduced (gloo here; ncclAllReduce inside the C-ABI on the GPUs) and must equal the
    whole-alignment lengths of the reference.
  * replicates (configs[4]): lvb_b200.sharding.replicate_shard gives each rank its chains; the
    gathered results must equal the single-process run of all chains.
  * candidates (SURVEY.md 8(e) third row): every rank holds the whole alignment, scores its range of
    the NNI neighbourhood's candidates (sharding.candidate_range), the lengths are gathered, and the
    best candidate is committed by every rank.
"""
import ctypes as C
import json
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from lvb_b200 import sharding  # noqa: E402


def test_word_blocks_tile_the_row():
    for nwords in (1, 2, 3, 62, 625, 6250, 62500, 62501):
        for world in (1, 2, 3, 4, 8):
            if (nwords + 1) // 2 < world:
                continue
            blocks = [sharding.word_block(nwords, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and sum(c for _, c in blocks) == nwords
            for (b0, c0), (b1, _) in zip(blocks, blocks[1:]):
                assert b0 + c0 == b1 and b1 % 2 == 0 and c0 >= 1
    with pytest.raises(ValueError):
        sharding.word_block(2, 1, 2)


def test_chain_ranges_tile_the_replicates():
    for reps in (1, 5, 64):
        for world in (1, 2, 4, 8):
            rs = [sharding.chain_range(reps, r, world) for r in range(world)]
            assert sum(c for _, c in rs) == reps and rs[0][0] == 0
            assert max(c for _, c in rs) - min(c for _, c in rs) <= 1
    assert sharding.replicate_shard(1000, 64, 3, 8) == (1024, 8)


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _mock_api():
    """ctypes view of the mock device's C-ABI (the subset the site-block path needs)"""
    import run_host_search
    from lvb_b200.gpu import _Topo

    L = C.CDLL(run_host_search.build_mock())
    vp, i64 = C.c_void_p, C.c_int64
    L.lvb_gpu_last_error.restype = C.c_char_p
    L.lvb_gpu_init.argtypes = [C.c_int, C.POINTER(vp)]
    L.lvb_gpu_matrix_upload_flat.argtypes = [vp, i64, i64, vp, i64, i64, i64]
    L.lvb_gpu_tree_create.argtypes = [vp, C.POINTER(C.c_int)]
    L.lvb_gpu_tree_commit.argtypes = [vp, C.c_int, C.POINTER(_Topo), C.POINTER(i64)]
    L.lvb_gpu_score_batch.argtypes = [vp, i64, C.POINTER(C.c_int), C.POINTER(_Topo), C.POINTER(i64)]
    L.lvb_gpu_tree_changes.argtypes = [vp, C.c_int, vp]
    L.lvb_gpu_nni_neighbourhood.argtypes = [vp, C.c_int, C.c_int, C.POINTER(_Topo), C.POINTER(i64)]
    L.lvb_gpu_shutdown.argtypes = [vp]
    return L


def _siteblock_worker(rank: int, world: int, port: int, out_path: str) -> None:
    import torch
    import torch.distributed as dist
    from lvb_b200.gpu import Topology

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        g = np.load(os.path.join(HERE, "golden", "example.npz"))
        enc = np.ascontiguousarray(g["enc"], dtype=np.uint64)
        n, nwords = enc.shape
        L = _mock_api()
        ctx = C.c_void_p()
        assert L.lvb_gpu_init(0, C.byref(ctx)) == 0
        w0, wc = sharding.word_block(nwords, rank, world)
        assert L.lvb_gpu_matrix_upload_flat(ctx, n, nwords, enc.ctypes.data, w0, wc, 2) == 0, L.lvb_gpu_last_error()
        t = C.c_int(-1)
        assert L.lvb_gpu_tree_create(ctx, C.byref(t)) == 0
        full = Topology(g["random_left"], g["random_right"], int(g["random_root"]))
        part = C.c_int64(0)
        ts = full.as_struct()
        assert L.lvb_gpu_tree_commit(ctx, t.value, C.byref(ts), C.byref(part)) == 0, L.lvb_gpu_last_error()
        lens = torch.tensor([part.value], dtype=torch.int64)
        dist.all_reduce(lens)  # the exchange step of the site-block path: K int64 partial lengths
        chg = np.zeros(2 * n - 3, dtype=np.int64)
        assert L.lvb_gpu_tree_changes(ctx, t.value, chg.ctypes.data) == 0
        chg_t = torch.from_numpy(chg.copy())
        dist.all_reduce(chg_t)
        # proposals scored on the blocks: the recorded NNI and the unchanged tree
        dirty = np.nonzero(g["nni_dirty"])[0]
        cands = [Topology(g["nni_left"], g["nni_right"], int(g["nni_root"]), dirty), Topology(full.left, full.right, full.root, [])]
        from lvb_b200.gpu import _Topo

        arr = (_Topo * 2)()
        for i, cnd in enumerate(cands):
            cnd.fill(arr[i])
        out = (C.c_int64 * 2)()
        trees = (C.c_int * 2)(t.value, t.value)
        assert L.lvb_gpu_score_batch(ctx, 2, trees, arr, out) == 0, L.lvb_gpu_last_error()
        sc = torch.tensor(list(out), dtype=torch.int64)
        dist.all_reduce(sc)
        # the exhaustive NNI neighbourhood on the blocks: 2(n-3) partial lengths, summed over the ranks
        s2 = C.c_int(-1)
        assert L.lvb_gpu_tree_create(ctx, C.byref(s2)) == 0
        nn = (C.c_int64 * (2 * (n - 3)))()
        assert L.lvb_gpu_nni_neighbourhood(ctx, t.value, s2.value, C.byref(ts), nn) == 0, L.lvb_gpu_last_error()
        nn_t = torch.tensor(list(nn), dtype=torch.int64)
        dist.all_reduce(nn_t)
        L.lvb_gpu_shutdown(ctx)
        if rank == 0:
            json.dump({"block": [w0, wc], "partial": int(part.value), "length": int(lens[0]), "scores": sc.tolist(),
                       "changes": chg_t.tolist(), "neighbourhood": nn_t.tolist()}, open(out_path, "w"))
    finally:
        dist.destroy_process_group()


def test_site_blocks_two_ranks_gloo(tmp_path):
    import torch.multiprocessing as mp

    import oracles

    out = str(tmp_path / "siteblock.json")
    mp.spawn(_siteblock_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = json.load(open(out))
    g = np.load(os.path.join(HERE, "golden", "example.npz"))
    want, _, chg = oracles.oracle_getplen(g["enc"], g["random_left"], g["random_right"], int(g["random_root"]))
    assert got["length"] == want == 11808
    assert 0 < got["partial"] < want                # rank 0 really held a part only
    assert got["scores"] == [11834, 11808]
    # per-node counts: every node's changes are the sum over blocks; only the root step is not cached
    assert got["changes"][g["enc"].shape[0]:] == [int(x) for x in chg[g["enc"].shape[0]:]]
    # neighbourhood lengths over blocks add up to whole-alignment lengths (sample against the oracle)
    from test_schedule import _nni_candidates

    n = g["enc"].shape[0]
    cands = list(_nni_candidates(g["random_left"], g["random_right"], int(g["random_root"]), n))
    assert len(got["neighbourhood"]) == len(cands) == 2 * (n - 3)
    for k in range(0, len(cands), 23):
        u, v, a, b, c, l2, r2 = cands[k]
        assert got["neighbourhood"][k] == oracles.oracle_getplen(g["enc"], l2, r2, int(g["random_root"]))[0], k


def _replicate_worker(rank: int, world: int, port: int, out_path: str, seed: int, reps: int) -> None:
    import torch.distributed as dist

    import run_host_search

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        C.CDLL(run_host_search.build_mock(), mode=C.RTLD_GLOBAL)  # the mock answers lvb_gpu_* for the host library
        from lvb_b200 import host

        g = np.load(os.path.join(HERE, "golden", "treelength.npz"))
        msa = host.HostMSA.from_rows([bytes(r) for r in g["tl5_thread_2_text"]])
        seed0, count = sharding.replicate_shard(seed, reps, rank, world)
        first, _ = sharding.chain_range(reps, rank, world)
        mine = []
        if count:
            res = msa.search(seed0, algorithm=1, replicates=count, host_threads=1)
            mine = [(first + i, r["final_length"], r["rearrangements"], r["topologies"], r["start_temperature"])
                    for i, r in enumerate(res)]
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)  # results only; nothing is exchanged on the data path
        if rank == 0:
            json.dump(sorted(x for part in everyone for x in part), open(out_path, "w"))
    finally:
        dist.destroy_process_group()


def test_replicates_sharded_over_two_ranks_gloo(tmp_path):
    """3 chains over 2 ranks (1 + 2) == the same 3 chains in one process"""
    import subprocess

    import torch.multiprocessing as mp

    out = str(tmp_path / "replicates.json")
    mp.spawn(_replicate_worker, args=(2, _free_port(), out, 1, 3), nprocs=2, join=True)
    got = json.load(open(out))
    p = subprocess.run([sys.executable, os.path.join(HERE, "run_host_search.py"), "--mock", "treelength.npz:tl5_thread_2_text",
                        "1", "1", "0", "3"], capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-2000:]
    one = json.loads(p.stdout.strip().splitlines()[-1])["results"]
    assert [x[0] for x in got] == [0, 1, 2]
    for r in range(3):
        assert got[r][1:] == [one[r]["final_length"], one[r]["rearrangements"], one[r]["topologies"], one[r]["start_temperature"]], r
    assert got[0][1] == 1846  # test_treelength_5_thread_2's pin is for seed 509739986; seed 1 reaches the same optimum


def _candidate_worker(rank: int, world: int, port: int, out_path: str) -> None:
    import torch
    import torch.distributed as dist
    from lvb_b200.gpu import Topology, _Topo
    from test_schedule import _nni_candidates

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        g = np.load(os.path.join(HERE, "golden", "example.npz"))
        enc = np.ascontiguousarray(g["enc"], dtype=np.uint64)
        n, nwords = enc.shape
        left, right, root = g["random_left"], g["random_right"], int(g["random_root"])
        L = _mock_api()
        ctx = C.c_void_p()
        assert L.lvb_gpu_init(0, C.byref(ctx)) == 0
        assert L.lvb_gpu_matrix_upload_flat(ctx, n, nwords, enc.ctypes.data, 0, nwords, 1) == 0  # the WHOLE alignment
        t = C.c_int(-1)
        assert L.lvb_gpu_tree_create(ctx, C.byref(t)) == 0
        ts = Topology(left, right, root).as_struct()
        cur = C.c_int64(0)
        assert L.lvb_gpu_tree_commit(ctx, t.value, C.byref(ts), C.byref(cur)) == 0, L.lvb_gpu_last_error()
        B = 2 * n - 3
        parent = np.full(B, -1, dtype=np.int64)
        for v in list(range(n, B)) + [root]:
            parent[left[v]] = v
            parent[right[v]] = v
        cands = []
        for u, v, a, b, c, l2, r2 in _nni_candidates(left, right, root, n):
            dirty, w = [], u
            while w != root:
                dirty.append(w)
                w = int(parent[w])
            cands.append(Topology(l2, r2, root, dirty))
        K = len(cands)
        first, count = sharding.candidate_range(K, rank, world)
        arr = (_Topo * count)()
        for i in range(count):
            cands[first + i].fill(arr[i])
        out = (C.c_int64 * count)()
        trees = (C.c_int * count)(*([t.value] * count))
        assert L.lvb_gpu_score_batch(ctx, count, trees, arr, out) == 0, L.lvb_gpu_last_error()
        # the exchange step: K lengths gathered (ranks hold ranges of different sizes: pad to the largest)
        width = (K + world - 1) // world
        mine = torch.full((width,), -1, dtype=torch.int64)
        mine[:count] = torch.tensor(list(out), dtype=torch.int64)
        parts = [torch.empty(width, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(parts, mine)
        lens = []
        for r in range(world):
            lens += parts[r][: sharding.candidate_range(K, r, world)[1]].tolist()
        assert len(lens) == K and min(lens) > 0
        # the winner (first of the shortest) is committed by EVERY rank on its own copy of the tree
        best = int(np.argmin(lens))
        bs = cands[best].as_struct()
        got = C.c_int64(0)
        assert L.lvb_gpu_tree_commit(ctx, t.value, C.byref(bs), C.byref(got)) == 0, L.lvb_gpu_last_error()
        agree = torch.tensor([got.value, best], dtype=torch.int64)
        both = [torch.empty(2, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(both, agree)
        L.lvb_gpu_shutdown(ctx)
        if rank == 0:
            json.dump({"current": int(cur.value), "lengths": lens, "best": best, "committed": [b.tolist() for b in both],
                       "range0": [first, count]}, open(out_path, "w"))
    finally:
        dist.destroy_process_group()


def test_candidates_sharded_over_two_ranks_gloo(tmp_path):
    import torch.multiprocessing as mp

    import oracles
    from test_schedule import _nni_candidates

    for world in (1, 2, 3, 8):
        rs = [sharding.candidate_range(194, r, world) for r in range(world)]
        assert rs[0][0] == 0 and sum(c for _, c in rs) == 194 and all(b0 + c0 == b1 for (b0, c0), (b1, _) in zip(rs, rs[1:]))
    out = str(tmp_path / "candidates.json")
    mp.spawn(_candidate_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = json.load(open(out))
    g = np.load(os.path.join(HERE, "golden", "example.npz"))
    enc, root = g["enc"], int(g["random_root"])
    n = enc.shape[0]
    cands = list(_nni_candidates(g["random_left"], g["random_right"], root, n))
    assert got["current"] == 11808 and len(got["lengths"]) == len(cands) == 2 * (n - 3)
    assert got["range0"] == [0, n - 3]  # rank 0 scored the first half only
    for k in list(range(0, len(cands), 17)) + [got["best"], len(cands) - 1]:
        u, v, a, b, c, l2, r2 = cands[k]
        assert got["lengths"][k] == oracles.oracle_getplen(enc, l2, r2, root)[0], k
    assert got["lengths"][got["best"]] == min(got["lengths"])
    assert got["committed"] == [[got["lengths"][got["best"]], got["best"]]] * 2  # both ranks hold the same new tree
