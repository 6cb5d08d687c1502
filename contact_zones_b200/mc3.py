"""K9 -- MC3 over several GPUs: chain sharding, the per-swap exchange of (log-likelihood, log-prior) and the snapshot
hand-off of the logged chain.

Reference semantics (`mcmc_generative.py:242-280`, quirk Q4): all chains run at temperature 1; a swap only exchanges
which chain is LABEL 0 (the logged one): `chain_swaps` times pick j in [1, n) and swap labels 0 <-> j with probability
min(1, exp((ll_j + prior_j) - (ll_0 + prior_0))).  No state moves between chains, so across GPUs the swap needs two
doubles per chain (one all-gather) and every rank then takes the same decisions from a shared counter-based stream.

One process per GPU (`torch.distributed`, backend nccl on the GPU box, gloo in the CPU tests); without an initialised
process group everything degenerates to the single-process case.
"""
import numpy as np


def _dist():
    try:
        import torch.distributed as dist
    except Exception:           # torch is plumbing only; single-process use does not need it
        return None
    return dist if (dist.is_available() and dist.is_initialized()) else None


class ChainShard:
    """Block partition of `n_chains` global chains over the ranks of the default process group."""

    def __init__(self, n_chains, rank=None, world=None):
        dist = _dist()
        self.world = int(world if world is not None else (dist.get_world_size() if dist else 1))
        self.rank = int(rank if rank is not None else (dist.get_rank() if dist else 0))
        self.n_chains = int(n_chains)
        base, extra = divmod(self.n_chains, self.world)
        self.counts = [base + (1 if r < extra else 0) for r in range(self.world)]
        self.offsets = np.concatenate([[0], np.cumsum(self.counts)]).astype(int)
        self.offset = int(self.offsets[self.rank])
        self.n_local = int(self.counts[self.rank])

    def owner(self, chain):
        """(rank, local index) of a global chain id."""
        r = int(np.searchsorted(self.offsets, chain, side='right') - 1)
        return r, int(chain - self.offsets[r])

    def is_local(self, chain):
        return self.owner(chain)[0] == self.rank

    # -- collectives ------------------------------------------------------------------------------------------
    def all_gather_scalars(self, ll_local, prior_local, device=None):
        """(ll, prior) of every global chain on every rank: 16 B per chain per swap period."""
        ll_local, prior_local = np.asarray(ll_local, float), np.asarray(prior_local, float)
        dist = _dist()
        if dist is None or self.world == 1:
            return ll_local.copy(), prior_local.copy()
        import torch
        m = max(self.counts)
        buf = torch.zeros(2 * m, dtype=torch.float64)
        buf[:self.n_local] = torch.from_numpy(ll_local)
        buf[m:m + self.n_local] = torch.from_numpy(prior_local)
        if device is not None:
            buf = buf.to(device)
        out = torch.empty(self.world * 2 * m, dtype=torch.float64, device=buf.device)
        dist.all_gather_into_tensor(out, buf)
        out = out.cpu().numpy().reshape(self.world, 2, m)
        ll = np.concatenate([out[r, 0, :self.counts[r]] for r in range(self.world)])
        pr = np.concatenate([out[r, 1, :self.counts[r]] for r in range(self.world)])
        return ll, pr

    def all_gather_device(self, ll_dev, prior_dev, out_row, scratch):
        """The same exchange without a host hop: `ll_dev` / `prior_dev` are device tensors (the context's own running
        arrays, `Engine.device_scalars`), `scratch` a persistent (2 * max_local) device buffer and `out_row` a persistent
        (world * 2 * max_local) device buffer -- two tiny device copies and ONE all-gather on the current stream;
        `unpack_gathered` turns rows read back later into per-chain arrays."""
        m = max(self.counts)
        scratch[:self.n_local].copy_(ll_dev)
        scratch[m:m + self.n_local].copy_(prior_dev)
        dist = _dist()
        if dist is None or self.world == 1:
            out_row.copy_(scratch)
        else:
            dist.all_gather_into_tensor(out_row, scratch)

    def unpack_gathered(self, rows):
        """rows: (..., world * 2 * max_local) host array -> (ll, prior) of shape (..., n_chains)."""
        rows = np.asarray(rows)
        m = max(self.counts)
        r = rows.reshape(rows.shape[:-1] + (self.world, 2, m))
        ll = np.concatenate([r[..., k, 0, :self.counts[k]] for k in range(self.world)], axis=-1)
        pr = np.concatenate([r[..., k, 1, :self.counts[k]] for k in range(self.world)], axis=-1)
        return ll, pr

    def all_gather_events(self, ll_local, prior_local, device=None):
        """The batched form: (n_events, n_local) arrays of every rank -> (n_events, n_chains), ONE collective for all
        the events of a persistent launch (16 B per chain and event)."""
        ll_local, prior_local = np.asarray(ll_local, float), np.asarray(prior_local, float)
        dist = _dist()
        if dist is None or self.world == 1:
            return ll_local.copy(), prior_local.copy()
        import torch
        e, m = ll_local.shape[0], max(self.counts)
        buf = np.zeros((2, e, m))
        buf[0, :, :self.n_local] = ll_local
        buf[1, :, :self.n_local] = prior_local
        t = torch.from_numpy(buf.ravel())
        if device is not None:
            t = t.to(device)
        out = torch.empty(self.world * t.numel(), dtype=torch.float64, device=t.device)
        dist.all_gather_into_tensor(out, t)
        out = out.cpu().numpy().reshape(self.world, 2, e, m)
        ll = np.concatenate([out[r, 0, :, :self.counts[r]] for r in range(self.world)], axis=1)
        pr = np.concatenate([out[r, 1, :, :self.counts[r]] for r in range(self.world)], axis=1)
        return ll, pr

    def broadcast_state(self, chain, arrays_local, template, device=None):
        """Every rank gets the state arrays of global `chain` from its owner (one broadcast of the packed arrays:
        ~0.4 MB at shape D, only at the N_SAMPLES logging points).  `arrays_local`: callable evaluated on the owner
        only, returns {name: array}; `template`: {name: (shape, dtype)}, identical on all ranks."""
        owner, _ = self.owner(chain)
        dist = _dist()
        if dist is None or self.world == 1:
            return arrays_local()
        import torch
        keys = sorted(template)
        n = int(sum(int(np.prod(template[k][0])) for k in keys))
        if self.rank == owner:
            arrs = arrays_local()
            flat = np.concatenate([np.asarray(arrs[k], dtype=np.float64).ravel() for k in keys])
            t = torch.from_numpy(np.ascontiguousarray(flat))
        else:
            t = torch.empty(n, dtype=torch.float64)
        if device is not None:
            t = t.to(device)
        dist.broadcast(t, src=owner)
        flat = t.cpu().numpy()
        out, o = {}, 0
        for k in keys:
            shape, dt = template[k]
            sz = int(np.prod(shape))
            out[k] = flat[o:o + sz].reshape(shape).astype(dt)
            o += sz
        return out


    def gather_states(self, requests, arrays_local, template, device=None):
        """Every rank gets the state arrays of a LIST of (global chain, key) requests, each from the chain's owner, in ONE
        collective: every rank packs the requests it owns into a matrix of (most requests of one owner, n) rows, the
        matrices are all-gathered (the collective the swap exchange uses anyway), and request r is row (position of r
        among its owner's requests) of its owner's matrix.  This is the per-launch form of `broadcast_state`: the logged
        snapshots of a whole launch move in one collective instead of one broadcast per logged sample.
        `arrays_local(key)`: evaluated on the owner only, returns {name: array}; `template` as in `broadcast_state`."""
        dist = _dist()
        if dist is None or self.world == 1:
            return [arrays_local(key) for _, key in requests]
        if not requests:
            return []
        import torch
        keys = sorted(template)
        n = int(sum(int(np.prod(template[k][0])) for k in keys))
        owners = [self.owner(chain)[0] for chain, _ in requests]
        pos, per_owner = [], [0] * self.world
        for o in owners:
            pos.append(per_owner[o])
            per_owner[o] += 1
        m = max(per_owner)
        rows = np.zeros((m, n))
        for r, (chain, key) in enumerate(requests):
            if owners[r] == self.rank:
                arrs = arrays_local(key)
                o = 0
                for k in keys:
                    sz = int(np.prod(template[k][0]))
                    rows[pos[r], o:o + sz] = np.asarray(arrs[k], dtype=np.float64).ravel()
                    o += sz
        t = torch.from_numpy(rows)
        if device is not None:
            t = t.to(device)
        allm = torch.empty(self.world * m * n, dtype=torch.float64, device=t.device)
        dist.all_gather_into_tensor(allm, t.reshape(-1))
        index = torch.tensor([o * m + p for o, p in zip(owners, pos)], device=t.device)
        flat = allm.view(self.world * m, n).index_select(0, index).cpu().numpy()
        out = []
        for r in range(len(requests)):
            d, o = {}, 0
            for k in keys:
                shape, dt = template[k]
                sz = int(np.prod(shape))
                d[k] = flat[r, o:o + sz].reshape(shape).astype(dt)
                o += sz
            out.append(d)
        return out


    def gather_records(self, requests, ring):
        """The device form of `gather_states`: `requests` = [(global chain, record index in the OWNER's ring)], `ring` =
        this rank's snapshot ring as a (n_records, record_bytes) uint8 device tensor (Engine.snapshot_ring).  Every rank
        selects the records it owns on the device, the ranks' (most records of one owner, record_bytes) blocks are
        all-gathered and ONE device-to-host copy returns the (n_requests, record_bytes) records in request order: no host
        hop on the owner, no per-sample transfer."""
        import torch
        dist = _dist()
        if not requests:
            return np.zeros((0, 0), np.uint8)
        owners = [self.owner(chain)[0] for chain, _ in requests]
        if dist is None or self.world == 1:
            idx = torch.tensor([r for _, r in requests], device=ring.device)
            return ring.index_select(0, idx).cpu().numpy()
        pos, per_owner = [], [0] * self.world
        for o in owners:
            pos.append(per_owner[o])
            per_owner[o] += 1
        m = max(per_owner)
        mine = [r for (_, r), o in zip(requests, owners) if o == self.rank]
        block = torch.zeros(m, ring.shape[1], dtype=torch.uint8, device=ring.device)
        if mine:
            block[:len(mine)] = ring.index_select(0, torch.tensor(mine, device=ring.device))
        allb = torch.empty(self.world * m, ring.shape[1], dtype=torch.uint8, device=ring.device)
        dist.all_gather_into_tensor(allb, block)
        index = torch.tensor([o * m + p for o, p in zip(owners, pos)], device=ring.device)
        return allb.index_select(0, index).cpu().numpy()


def swap_chains(chain_idx, ll, prior, chain_swaps, rng, statistics=None):
    """`MCMCGenerative.swap_chains` (mcmc_generative.py:242-280) on gathered scalars.  `chain_idx` (list, label ->
    chain) is permuted in place; `rng` is a numpy Generator every rank seeds identically, so all ranks agree."""
    n = len(chain_idx)
    for _ in range(int(chain_swaps)):
        if statistics is not None:
            statistics['n_swaps'] += 1
        swap_from = chain_idx[0]
        swap_to_idx = int(rng.integers(1, n))                 # np.random.choice(range(1, n_chains))   :252
        swap_to = chain_idx[swap_to_idx]
        mh = (ll[swap_to] - ll[swap_from]) + (prior[swap_to] - prior[swap_from])     # :263-265 with q_to = q_from = 0
        if np.log(rng.random()) < mh:                         # :268
            chain_idx[0], chain_idx[swap_to_idx] = swap_to, swap_from
            if statistics is not None:
                statistics['accepted_swaps'] += 1
    return chain_idx
