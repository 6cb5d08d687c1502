"""CPU, world_size 2 over gloo: the multi-GPU plumbing of the MC3 swap (K9, contact_zones_b200/mc3.py) -- chain
sharding, the all-gather of (ll, prior), identical swap decisions on every rank, the snapshot broadcast of the logged
chain -- against the single-process result."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_CHAINS = 7          # uneven split: 4 + 3


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _chain_values(seed=5):
    rng = np.random.default_rng(seed)
    return rng.normal(-1000., 3., N_CHAINS), rng.normal(-20., 1., N_CHAINS)


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from contact_zones_b200.mc3 import ChainShard, swap_chains
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    shard = ChainShard(N_CHAINS)
    ll, pr = _chain_values()
    lo, hi = shard.offset, shard.offset + shard.n_local
    g_ll, g_pr = shard.all_gather_scalars(ll[lo:hi], pr[lo:hi])
    chain_idx = list(range(N_CHAINS))
    stats = {'n_swaps': 0, 'accepted_swaps': 0}
    rng = np.random.default_rng([11, 0])
    for _ in range(20):
        swap_chains(chain_idx, g_ll, g_pr, 3, rng, stats)
    # snapshot of the logged chain: only its owner evaluates the callable
    calls = []

    def local_arrays():
        calls.append(rank)
        c = chain_idx[0]
        return {'zones': np.arange(12).reshape(3, 4) % 2 == (c % 2), 'weights': np.full((2, 3), float(c))}

    snap = shard.broadcast_state(chain_idx[0], local_arrays, {'zones': ((3, 4), bool), 'weights': ((2, 3), float)})
    # the per-launch form: several (chain, key) requests in one collective, every owner evaluated only for its own
    asked = []

    def arrays_of(key):
        asked.append(key)
        return {'zones': np.arange(12).reshape(3, 4) % 3 == (key % 3), 'weights': np.full((2, 3), -0.5 * key)}

    reqs = [(c, 10 * c + i) for i, c in enumerate([0, 6, 3, 4, 6])]
    many = shard.gather_states(reqs, arrays_of, {'zones': ((3, 4), bool), 'weights': ((2, 3), float)})
    # ... and its device form on raw ring records: rank r's ring holds 100 r + record index in every byte
    import torch
    ring = (torch.arange(6, dtype=torch.uint8)[:, None] + 100 * rank).repeat(1, 32)
    rec_reqs = [(0, 2), (6, 1), (5, 5), (3, 0), (4, 3)]                     # (global chain, record index on its owner)
    recs = shard.gather_records(rec_reqs, ring)
    out[rank] = dict(counts=shard.counts, offset=shard.offset, ll=g_ll, pr=g_pr, idx=chain_idx, stats=stats,
                     owner=shard.owner(chain_idx[0])[0], calls=calls, w=snap['weights'], z=snap['zones'],
                     asked=asked, many=[(m['zones'], m['weights']) for m in many], reqs=reqs, recs=recs, rec_reqs=rec_reqs)
    dist.destroy_process_group()


def test_mc3_exchange_world2():
    from contact_zones_b200.mc3 import ChainShard, swap_chains
    port = _free_port()
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    r0, r1 = out[0], out[1]
    ll, pr = _chain_values()
    assert r0['counts'] == [4, 3] and (r0['offset'], r1['offset']) == (0, 4)
    for r in (r0, r1):
        assert np.array_equal(r['ll'], ll) and np.array_equal(r['pr'], pr)
    # single-process reference of the same swap sequence
    idx = list(range(N_CHAINS))
    stats = {'n_swaps': 0, 'accepted_swaps': 0}
    rng = np.random.default_rng([11, 0])
    for _ in range(20):
        swap_chains(idx, ll, pr, 3, rng, stats)
    assert r0['idx'] == r1['idx'] == idx and sorted(idx) == list(range(N_CHAINS))
    assert r0['stats'] == r1['stats'] == stats and stats['n_swaps'] == 60 and 0 < stats['accepted_swaps'] < 60
    owner = r0['owner']
    assert (out[owner]['calls'], out[1 - owner]['calls']) == ([owner], [])
    assert np.array_equal(r0['w'], r1['w']) and r0['w'][0, 0] == float(idx[0]) and np.array_equal(r0['z'], r1['z'])
    s = ChainShard(N_CHAINS, rank=0, world=1)
    assert s.n_local == N_CHAINS and s.owner(5) == (0, 5)

    # gather_states: identical results on both ranks, each request evaluated on its owner only (chains 0-3 -> rank 0)
    assert r0['asked'] == [k for c, k in r0['reqs'] if c < 4] and r1['asked'] == [k for c, k in r0['reqs'] if c >= 4]
    for (c, key), (z0, w0), (z1, w1) in zip(r0['reqs'], r0['many'], r1['many']):
        assert np.array_equal(z0, np.arange(12).reshape(3, 4) % 3 == (key % 3)) and z0.dtype == bool
        assert np.array_equal(w0, np.full((2, 3), -0.5 * key))
        assert np.array_equal(z0, z1) and np.array_equal(w0, w1)
    # gather_records: request (chain, i) returns record i of the chain's owner, on both ranks
    want = np.array([[100 * (c >= 4) + i] * 32 for c, i in r0['rec_reqs']], np.uint8)
    assert np.array_equal(r0['recs'], want) and np.array_equal(r1['recs'], want)

def test_swap_accepts_better_chain():
    """mcmc_generative.py:263-274: label 0 moves to a chain whose posterior is higher with probability 1."""
    from contact_zones_b200.mc3 import swap_chains
    idx = [0, 1]
    swap_chains(idx, np.array([-10., -5.]), np.array([0., 0.]), 1, np.random.default_rng(0))
    assert idx == [1, 0]
    n_acc = 0
    for s in range(400):
        idx = [0, 1]
        swap_chains(idx, np.array([-5., -6.]), np.array([0., 0.]), 1, np.random.default_rng(s))
        n_acc += idx == [1, 0]
    assert abs(n_acc / 400 - np.exp(-1.)) < 0.08
