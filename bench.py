"""bench.py -- the driver's measurement contract for the B200-native sBayes MCMC inner loop.

    python bench.py --gpus N --steps K --warmup W                 (N>1: launched by torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W  (CPU arm: the oracle port on the host cores)

Metric (BASELINE.json): MCMC proposals/s over batched MC3 chains; likelihood GB/s vs the HBM peak.
Workload: configs[3] "synthetic 2000 objects x 200 features x 8 states, 7 zones, 20 families, 64 MC3 chains":
64 chains IN TOTAL, sharded over the N GPUs (strong scaling, the configuration north_star names; the chains are
independent Markov chains, sharded with no data-path collective; the only exchange is the MC3 swap's all-gather of
(ll, prior), 16 B per chain, once per step).  For N > 1 the line also carries `weak`: 64 chains resident PER GPU.

One bench "step" = `--mcmc-steps` proposals on every chain in one persistent-kernel launch (K8) followed by the
MC3 swap exchange (mcmc_generative.py:242-280) -- the unit the reference's `generate_samples` repeats.
  value      proposals/s with every chain state resident in HBM (device-timed, CUDA events, max over ranks); chains
             initialised by the product's generate_initial_sample path and warmed; median over seeds 0 / 1 / 2
  e2e        the same through the C-ABI with HOST buffers: every step uploads all chain states from pinned host
             memory (sb_set_states), runs the same launch and reads the logged chain's state + all (ll, prior) back
  api_e2e    ZoneMCMC.generate_samples at 1000 steps per logged sample (the drop-in API, snapshot ring inside)
  roofline   K1 (full mixture log-likelihood, lh_full_kernel) on configs[4]'s largest shape (256 chains x 20000 x
             100 x 4), timed live with CUDA events on the launching stream; algorithmic bytes per SURVEY.md 8(d)
  cpu_baseline  the oracle port of the reference sampler (oracle/sbayes_sampler_oracle.py) on the host cores
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PRECISION = {'weights': 15, 'universal': 40, 'contact': 20, 'inheritance': 20}     # default_config.json:6-11
D_SHAPE = dict(N=2000, F=200, S=8, Z=7, Fam=20)
E_SHAPE = dict(N=20000, F=100, S=4, Z=4, zone_size=25, chains=256)
METRIC = 'mcmc_proposals_per_sec'


def d_data():
    from contact_zones_b200 import synthetic
    return synthetic.make_data(D_SHAPE['N'], D_SHAPE['F'], D_SHAPE['S'], n_families=D_SHAPE['Fam'], seed=0,
                               with_dist=False, planted_zones=D_SHAPE['Z'], planted_size=20)


def d_state(data, seed):
    from contact_zones_b200 import synthetic
    return synthetic.random_state(data, D_SHAPE['Z'], 5, True, True, seed=seed)


def default_ops():
    """mcmc_setup.py:70-95 with the default STEPS (default_config.json:13-20), SAMPLE_SOURCE + INHERITANCE."""
    return {'shrink_zone': .02, 'grow_zone': .02, 'swap_zone': .01, 'gibbs_sample_weights': .4,
            'gibbs_sample_p_global': .05, 'gibbs_sample_p_zones': .4, 'gibbs_sample_p_families': .1}


def b_lh(N, F, S, Z, Fam, C, src, inh):
    """algorithmic bytes of one full log-likelihood evaluation of one chain."""
    return N * F + src * N * F + N + inh * 2 * N + (1 + Z + Fam) * F * S * 8 + F * C * 8 + 8


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference sampler, one chain per process on every host core
# ---------------------------------------------------------------------------------------------------------
_CPU = {}


def _cpu_init(seed_base):
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    os.environ['OMP_NUM_THREADS'] = '1'
    import sbayes_sampler_oracle as so           # checker / CPU baseline only
    data = d_data()
    ident = mp.current_process()._identity
    wid = ident[0] if ident else 0
    st = d_state(data, seed_base + wid)
    cfg = {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
           'universal': {'type': 'uniform'}, 'contact': {'type': 'uniform'}, 'inheritance': {'type': 'uniform'}}
    _CPU['s'] = so.OracleSampler(data.features, data.states, data.families, data.network['adj_mat'], True, True, cfg,
                                 default_ops(), PRECISION, 3, 50, 0.85, st, seed=seed_base + wid)


def _cpu_run(n):
    t0 = time.perf_counter()
    _CPU['s'].run(n)
    return time.perf_counter() - t0


def cpu_pool(cores):
    ctx = mp.get_context('fork')
    pool = ctx.Pool(cores, initializer=_cpu_init, initargs=(1000,))
    pool.map(_cpu_run, [1] * cores)              # every worker built + one step done
    return pool


def cpu_measure(pool, cores, proposals_per_proc):
    t0 = time.perf_counter()
    pool.map(_cpu_run, [proposals_per_proc] * cores, chunksize=1)
    dt = time.perf_counter() - t0
    return cores * proposals_per_proc / dt, dt


def cpu_likelihood(n_eval=3):
    """One full likelihood evaluation (Likelihood.__call__(caching=False), model.py:145-184, as restated by the oracle)
    of ONE chain on the roofline shape E at N=20000, single thread -- SURVEY.md 8(d)."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import sbayes_oracle as orc                  # checker / CPU baseline only
    from contact_zones_b200 import synthetic
    e = E_SHAPE
    data = synthetic.make_data(e['N'], e['F'], e['S'], n_families=0, seed=2, with_dist=False)
    st = synthetic.random_state(data, e['Z'], e['zone_size'], False, False, seed=10)
    fam = np.zeros((0, e['N']), bool)
    best = float('inf')
    for _ in range(n_eval):
        t0 = time.perf_counter()
        orc.log_likelihood(data.features, fam, False, st['zones'], st['weights'], st['p_global'], st['p_zones'], None)
        best = min(best, time.perf_counter() - t0)
    nbytes = b_lh(e['N'], e['F'], e['S'], e['Z'], 0, 2, 0, 0)
    return {'likelihood_eval_ms': 1e3 * best, 'likelihood_gbs': nbytes / best / 1e9,
            'likelihood_sample': f'1 chain x {e["N"]} x {e["F"]} x {e["S"]}, 1 thread, best of {n_eval}'}


def cpu_name():
    try:
        for line in open('/proc/cpuinfo'):
            if line.startswith('model name'):
                return line.split(':', 1)[1].strip()
    except OSError:
        pass
    return 'unknown'


def run_reference_arm(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    pool = cpu_pool(cores)
    per = args.cpu_proposals
    for _ in range(args.warmup):
        cpu_measure(pool, cores, max(1, per // 4))
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_measure(pool, cores, per)
    dt = time.perf_counter() - t0
    pool.close()
    value = args.steps * cores * per / dt
    sample = f'{cores} processes x 1 chain x {per} proposals per step on shape D (oracle port of ZoneMCMC.step)'
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'proposals/s', 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True,
        'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(args, 1),
        'cpu_baseline': {'value': value, 'unit': 'proposals/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'cpu': cpu_name()},
        'e2e': {'value': value, 'unit': 'proposals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0}), flush=True)


def workload_config(args, world):
    """configs[3] of BASELINE.json.  The dict does not depend on the number of GPUs: the 64 chains are SHARDED over
    them (strong scaling), and the CPU arm prints the same dict."""
    return {'workload': 'synthetic 2000x200x8, 7 zones, 20 families, 64 MC3 chains sharded over the GPUs, SAMPLE_SOURCE, '
                        'default operator mix', 'chains_total': args.chains_total,
            'mcmc_steps_per_step': args.mcmc_steps, 'swap_exchange': 'all_gather(ll, prior) per step',
            'initial_state': 'generate_initial_sample (random connected zones of 5 sites, smoothed-MLE probabilities, '
                             'Gibbs-drawn sources) + %d untimed steps per chain' % args.chain_warm_steps,
            'seeds': list(args.seeds), 'l2': 'L2 flushed (256 MiB write) between timed steps'}


# ---------------------------------------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML polled every 5 ms from a thread (the
    nvidia-smi loop of the profiling recipe cannot sample faster than ~100 ms; it stays as the fallback)."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
    BITS = {0x8: 'hw_slowdown', 0x40: 'hw_thermal_slowdown', 0x20: 'sw_thermal_slowdown', 0x4: 'sw_power_cap',
            0x80: 'hw_power_brake_slowdown'}

    def __init__(self, index):
        self.rows, self.proc, self.nvml, self.stop = [], None, None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={index}', f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _poll(self):
        nv = self.nvml
        while not self.stop:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.rows.append((time.perf_counter(), mhz, mask))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(',')]))

    def summary(self, t0, t1):
        if self.nvml is not None:
            self.stop = True
            rows = [r for r in self.rows if t0 <= r[0] <= t1] or self.rows[-3:]
            mask = 0
            for r in rows:
                mask |= r[2]
            return {'sm_mhz': float(np.median([r[1] for r in rows])) if rows else None, 'sm_max_mhz': self.max_mhz,
                    'reasons': sorted(n for b, n in self.BITS.items() if mask & b), 'samples': len(rows),
                    'source': 'nvml, 5 ms period'}
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows[-3:]]
        sm = [float(r[0]) for r in rows if r[0].replace('.', '').isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace('.', '').isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = sorted({names[i] for r in rows for i in range(4) if len(r) > 2 + i and r[2 + i] == 'Active'})
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(rows), 'source': 'nvidia-smi -lms 100'}


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def k1_traffic(pattern='r*_k1_ncu_full_N20000.txt'):
    """DRAM bytes per launch of a kernel (default: K1 at N=20000) from the committed `ncu --set full` summary
    (profiles/), or None."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', pattern)))
    if not files:
        return None
    tot, unit = 0.0, {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    for line in open(files[-1]):
        m = re.match(r'dram__bytes_(read|write)\.sum\s+([0-9.,]+)\s+(\w+)', line)
        if m:
            tot += float(m.group(2).replace(',', '')) * unit.get(m.group(3), 1.0)
    return tot or None


def measured_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']), 'measured'
    except Exception:
        return 6650.0, 'fallback'


def likelihood_leg(torch, device, n_launch=60, sizes=(500, 1000, 2000, 5000, 10000, 20000)):
    """K1 on configs[4] (256 chains x N x 100 x 4, Z=4 zones of 25, C=2, no source): GB/s per SURVEY.md 8(d)."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    out = []
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0, 'the legacy default stream cannot be handed to the C ABI (0 selects the context\'s own stream)'
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    for N in sizes:
        e = E_SHAPE
        data = synthetic.make_data(N, e['F'], e['S'], n_families=0, seed=2, with_dist=False)
        eng = Engine(state_idx(data), data.states, None, data.network['adj_mat'], None, n_zones=e['Z'],
                     n_chains=e['chains'], inheritance=False, sample_source=False, device=device)
        eng.set_stream(stream)
        sts = [synthetic.random_state(data, e['Z'], min(e['zone_size'], N // 8), False, False, seed=10 + i)
               for i in range(8)]
        for c in range(e['chains']):
            s = sts[c % 8]
            eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], None, None)
        for _ in range(5):
            eng.eval_likelihood_async()
        torch.cuda.synchronize()
        # every timed launch runs behind an L2 flush (256 MiB write): the flush is enqueued first, so by the time the
        # GPU reaches the start event the launch is already queued and the event pair brackets the kernel alone
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_launch)]
        for a, b in evs:
            flush.fill_(1)
            a.record()
            eng.eval_likelihood_async()
            b.record()
        torch.cuda.synchronize()
        ms = np.array([a.elapsed_time(b) for a, b in evs])
        # the same launches back to back inside ONE event pair (warm L2, launch gaps included)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n_launch):
            eng.eval_likelihood_async()
        b.record()
        torch.cuda.synchronize()
        us_b2b = a.elapsed_time(b) * 1e3 / n_launch
        nbytes = e['chains'] * b_lh(N, e['F'], e['S'], e['Z'], 0, 2, 0, 0)
        ll = eng.read_ll()
        out.append({'N': N, 'chains': e['chains'], 'bytes_per_launch': nbytes, 'us_mean': float(ms.mean() * 1e3),
                    'us_min': float(ms.min() * 1e3), 'us_back_to_back': float(us_b2b),
                    'gbs_mean': float(nbytes / ms.mean() / 1e6),
                    'll_chain0': float(ll[0])})
        eng.close()
    return out


def source_free_leg(torch, args, data, states, device, stream, n_launch=3):
    """Shape D with SAMPLE_SOURCE = false: zone moves + alter_weights / alter_p_* (mcmc_setup.py:87-95), device-timed."""
    from contact_zones_b200.engine import Engine
    nc = len(states)
    eng = Engine(state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=D_SHAPE['Z'],
                 n_chains=nc, inheritance=True, sample_source=False, device=device)
    eng.set_stream(stream)
    for c, s in enumerate(states):
        eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], None)
    eng.eval()
    ops = {'shrink_zone': .02, 'grow_zone': .02, 'swap_zone': .01, 'alter_weights': .4, 'alter_p_global': .05,
           'alter_p_zones': .4, 'alter_p_families': .1}
    eng.set_operators(ops, PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=args.seed)
    eng.run(args.mcmc_steps)
    ms = []
    for _ in range(n_launch):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        eng.run(args.mcmc_steps)
        b.record()
        torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    ll_run, _ = eng.chain_scalars()
    ll_new, _, _ = eng.eval()
    acc, rej = eng.op_stats()
    eng.close()
    t = float(np.mean(ms))
    return {'value': nc * args.mcmc_steps / (t * 1e-3), 'unit': 'proposals/s', 'ms_per_launch': t, 'launches': n_launch,
            'us_per_proposal_per_chain': 1e3 * t / args.mcmc_steps, 'operators': ops,
            'accept_rate': float(acc.sum() / max(1, acc.sum() + rej.sum())),
            'll_drift_rel': float(np.max(np.abs(ll_run - ll_new) / np.abs(ll_new)))}


def state_idx(data):
    from contact_zones_b200.engine import state_index
    return state_index(data.features)


MODEL_CFG = {'N_AREAS': D_SHAPE['Z'], 'MIN_M': 3, 'MAX_M': 50, 'INHERITANCE': True, 'SAMPLE_SOURCE': True,
             'PRIOR': {'geo': {'type': 'uniform'}, 'area_size': {'type': 'none'}, 'weights': {'type': 'uniform'},
                       'universal': {'type': 'uniform'}, 'contact': {'type': 'uniform'},
                       'inheritance': {'type': 'uniform'}}}            # default_config.json:27-41


def make_sampler(data, n_chains, seed, mcmc_steps, model_cfg=None, ops=None):
    """The product's sampler object for `n_chains` MC3 chains (sharded over the ranks of the process group by
    ChainShard): initial states come from its own `generate_initial_sample` path (zone_sampling.py:1197-1233: random
    connected zones of M_INITIAL sites, uniform weights, smoothed-MLE probabilities, sources by a Gibbs sweep)."""
    from contact_zones_b200.model import Model
    from contact_zones_b200.sampling import ZoneMCMC
    model = Model(data, dict(model_cfg or MODEL_CFG))
    return ZoneMCMC(data=data, model=model, n_chains=n_chains, operators=ops or default_ops(), var_proposal=PRECISION,
                    p_grow_connected=0.85, initial_size=5, mc3=n_chains > 1, swap_period=mcmc_steps, chain_swaps=3,
                    seed=seed)


def sampler_leg(torch, args, data, n_chains, seed, stream, barrier, flush, timed_steps, warm_steps, keep_states=False):
    """Device-resident leg for one seed: chains initialised and sharded by the product (`ZoneMCMC`), warmed, then
    `timed_steps` bench steps of `--mcmc-steps` proposals per chain, each followed by the MC3 exchange -- an NCCL
    all-gather straight from the contexts' device (ll, prior) arrays into a persistent device buffer
    (`ChainShard.all_gather_device`); the swaps are replayed from ONE read-back after the timed region (labels only
    move, mcmc_generative.py:270-271)."""
    import numpy as np
    from contact_zones_b200.mc3 import swap_chains
    smp = make_sampler(data, n_chains, seed, args.mcmc_steps)
    eng = smp._initialise_chains()
    eng.sync()
    eng.set_stream(stream)
    shard = smp.shard
    m = max(shard.counts)
    ll_dev, pr_dev = eng.device_scalars()
    scratch = torch.zeros(2 * m, dtype=torch.float64, device='cuda')
    gathered = torch.zeros(timed_steps + warm_steps + 1, shard.world * 2 * m, dtype=torch.float64, device='cuda')
    eng.run(args.chain_warm_steps)                       # untimed: the chains leave their initial states
    for w in range(warm_steps):
        eng.run(args.mcmc_steps)
        shard.all_gather_device(ll_dev, pr_dev, gathered[w], scratch)
    barrier()
    acc0, rej0 = eng.op_stats()
    l0 = eng.kernel_launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(timed_steps)]
    t0 = time.perf_counter()
    for k, (a, b) in enumerate(ev):
        flush.fill_(1)
        a.record()
        eng.run(args.mcmc_steps)
        b.record()
        shard.all_gather_device(ll_dev, pr_dev, gathered[warm_steps + k], scratch)
    end = torch.cuda.Event(enable_timing=True)
    end.record()
    barrier()
    t1 = time.perf_counter()
    launches = eng.kernel_launches - l0
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    # a bench step = launch + MC3 exchange (+ the 40 us L2 flush in front of it): first launch start -> end of the last
    # exchange, on the device clock
    step_ms = max(ev[0][0].elapsed_time(end) / timed_steps, kernel_ms)
    exch_ms = step_ms - kernel_ms
    # replay of the swap decisions from the gathered scalars (host, after the timed region)
    rows = gathered[warm_steps:warm_steps + timed_steps].cpu().numpy()
    ll_all, pr_all = shard.unpack_gathered(rows)
    stats = {'n_swaps': 0, 'accepted_swaps': 0}
    for k in range(timed_steps):
        swap_chains(smp.chain_idx, ll_all[k], pr_all[k], smp.chain_swaps or 0, smp._swap_rng, stats)
    # drift of the incrementally maintained (ll, prior) against a from-scratch evaluation (K1 with source + K2)
    ll_run, pr_run = eng.chain_scalars()
    ll_new, pr_new, _ = eng.eval()
    drift = float(np.max(np.abs(ll_run - ll_new) / np.abs(ll_new)))
    assert np.allclose(ll_all[-1][shard.offset:shard.offset + shard.n_local], ll_run, rtol=0, atol=0)
    acc1, rej1 = eng.op_stats()
    acc, rej = (acc1 - acc0).sum(axis=0), (rej1 - rej0).sum(axis=0)
    states = [eng.get_state(c, one_hot_source=False) for c in range(shard.n_local)] if keep_states else None
    out = {'seed': seed, 'states': states, 'step_ms': step_ms, 'kernel_ms': kernel_ms, 'exchange_ms': exch_ms, 'launches': int(launches),
           'drift': drift, 'acc': acc, 'rej': rej, 'swaps': stats, 't0': t0, 't1': t1, 'n_local': shard.n_local,
           'team': int(os.environ.get('SBAYES_B200_TEAM', 0)) or None,
           'zone_size_mean': float(np.mean([eng.get_state(c, fields=('zones',))['zones'].sum() for c in range(min(4, shard.n_local))]))}
    smp.close()
    return out


def host_states(states):
    """The rank's chain states as stacked, pinned host arrays (the e2e leg's inputs): the states the device-resident leg
    left its chains in (`Engine.get_state` of every local chain after the timed steps), so that both legs time the same
    workload -- warmed chains -- and differ only in where the data lives."""
    import torch
    arrs = dict(zones=np.stack([s['zones'] for s in states]).astype(np.uint8), weights=np.stack([s['weights'] for s in states]),
                p_global=np.stack([s['p_global'][0] for s in states]), p_zones=np.stack([s['p_zones'] for s in states]),
                p_families=np.stack([s['p_families'] for s in states]),
                source=np.stack([s['source'] for s in states]).astype(np.uint8))
    return {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in arrs.items()}


def e2e_leg(torch, args, data, nc, chain_offset, device, seed, barrier, n_steps, states):
    """The same bench step through the C ABI with HOST buffers: every step uploads all chain states of the rank from
    pinned host memory (sb_set_states), runs the launch and reads the logged chain's state and every chain's
    (ll, prior) back.  Two contexts alternate, so the upload of step k+1 (copy engine + host packing) overlaps the
    launch of step k; every step's H2D and D2H are inside the timed region."""
    from contact_zones_b200.engine import Engine
    host = host_states(states)
    assert host['zones'].shape[0] == nc
    h2d = sum(t.numel() * t.element_size() for t in host.values())
    engs = []
    for i in range(2):
        e = Engine(state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=D_SHAPE['Z'],
                   n_chains=nc, inheritance=True, sample_source=True, device=device, chain_offset=chain_offset)
        engs.append(e)

    def upload(e):
        e.set_states(0, host['zones'].numpy(), host['weights'].numpy(), host['p_global'].numpy(),
                     host['p_zones'].numpy(), host['p_families'].numpy(), host['source'].numpy())

    def read(e):
        s0 = e.get_state(0, one_hot_source=False)
        ll, pr = e.chain_scalars()
        return sum(v.nbytes for v in s0.values() if v is not None) + ll.nbytes + pr.nbytes

    for e in engs:                  # untimed pass: operators configured, staging buffers warm
        upload(e)
        e.eval()
        e.set_operators(default_ops(), PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=seed)
        e.run(args.mcmc_steps)
        read(e)
    l0 = sum(e.kernel_launches for e in engs)
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for k in range(n_steps):
        e = engs[k & 1]
        upload(e)
        e.run(args.mcmc_steps)
        if k >= 1:
            d2h = read(engs[(k - 1) & 1])
    d2h = read(engs[(n_steps - 1) & 1])
    barrier()
    dt = (time.perf_counter() - t0) / n_steps
    launches = sum(e.kernel_launches for e in engs) - l0
    # the serial form (upload -> launch -> read back on one context) for the breakdown
    e = engs[0]
    t0 = time.perf_counter()
    for k in range(3):
        upload(e)
    e.sync()
    up_s = (time.perf_counter() - t0) / 3
    # ... and the whole step on ONE context (upload -> launch -> read back, nothing overlapped)
    t0 = time.perf_counter()
    for k in range(3):
        upload(e)
        e.run(args.mcmc_steps)
        read(e)
    serial_s = (time.perf_counter() - t0) / 3
    for e in engs:
        e.close()
    return {'s_per_step': dt, 'h2d': int(h2d), 'd2h': int(d2h), 'launches': int(launches), 'upload_ms': 1e3 * up_s,
            'upload_gbs': h2d / up_s / 1e9, 'serial_s_per_step': serial_s}


def api_leg(torch, data, n_chains, n_steps, n_samples, seed, model_cfg=None, ops=None):
    """The drop-in API end to end: `ZoneMCMC.generate_samples(N_STEPS, N_SAMPLES)` at the reference's default of 1000
    steps per logged sample (default_config.json:3-4), MC3 swaps every 1000 steps; the rate is taken from the
    `sampling_time` the sampler records itself (mcmc_generative.py:233-235: the loop, logging and swaps included)."""
    # one short untimed run first: CUDA module loading, NCCL channels and pinned staging are one-time costs of the
    # process, and the timed run below lasts only a few tenths of a second
    warm = make_sampler(data, n_chains, seed, 1000, model_cfg, ops)
    warm.generate_samples(3000, 3)
    warm.close()
    smp = make_sampler(data, n_chains, seed, 1000, model_cfg, ops)
    smp.generate_samples(n_steps, n_samples)
    st = smp.statistics
    out = {'value': n_chains * n_steps / st['sampling_time'], 'unit': 'proposals/s', 'n_chains': n_chains,
           'n_steps': n_steps, 'n_samples': n_samples, 'sampling_time_s': st['sampling_time'],
           'logged_samples': len(st['sample_zones']), 'swap_ratio': st['swap_ratio'],
           'us_per_proposal_per_chain': 1e6 * st['sampling_time'] / n_steps}
    smp.close()
    return out


def time_launches(torch, fn, flush, n_launch=40, warm=5):
    """Mean / min device time of `fn`'s launches, each behind an L2 flush that is enqueued first (the event pair then
    brackets the kernel alone: no host enqueue gap inside it, cold L2)."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_launch)]
    for a, b in evs:
        flush.fill_(1)
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ms = np.array([a.elapsed_time(b) for a, b in evs])
    return float(ms.mean() * 1e3), float(ms.min() * 1e3)


def shape_d_kernel_leg(torch, data, device, stream, peak):
    """Roofline fractions of the full-evaluation kernels on the headline shape D (64 chains x 2000 x 200 x 8, 7 zones,
    20 families): K1 (source-free likelihood), K1s (likelihood at the chosen source) and K4 (full recount of the
    count tables); algorithmic bytes per SURVEY.md 8(d), CUDA events on the launching stream."""
    from contact_zones_b200 import synthetic
    from contact_zones_b200.engine import Engine
    d = D_SHAPE
    nc = 64
    eng = Engine(state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=d['Z'],
                 n_chains=nc, inheritance=True, sample_source=True, device=device)
    eng.set_stream(stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    sts = [synthetic.random_state(data, d['Z'], 20, True, True, seed=300 + i) for i in range(8)]
    for c in range(nc):
        s = sts[c % 8]
        eng.set_state(c, s['zones'], s['weights'], s['p_global'], s['p_zones'], s['p_families'], s['source'])
    out = {}
    table_bytes = (1 + d['Z'] + d['Fam']) * d['F'] * d['S']
    for key, fn, per_chain in (
            ('k1', lambda: eng.eval_likelihood_async(use_source=False), b_lh(d['N'], d['F'], d['S'], d['Z'], d['Fam'], 3, 0, 1)),
            ('k1s', lambda: eng.eval_likelihood_async(use_source=True), b_lh(d['N'], d['F'], d['S'], d['Z'], d['Fam'], 3, 1, 1)),
            # K4 reads x, source, zone ids, family ids and writes the int32 count tables (+ the [F][4][3] source table)
            ('k4', lambda: eng.recount_async(), 2 * d['N'] * d['F'] + 3 * d['N'] + 4 * table_bytes + 48 * d['F'])):
        us, us_min = time_launches(torch, fn, flush)
        nbytes = nc * per_chain
        summary = {'k1': 'r*_k1_free_block_ncu_full_shapeD.txt', 'k1s': 'r*_k1s_source_block_ncu_full_shapeD.txt',
                   'k4': 'r*_k4_count_block_ncu_full_shapeD.txt'}[key]
        out[key] = {'bound': 'hbm', 'achieved': nbytes / us / 1e3, 'peak': peak, 'unit': 'GB/s',
                    'frac': nbytes / us / 1e3 / peak, 'traffic': k1_traffic(summary), 'bytes_per_launch': int(nbytes),
                    'kernel': {'k1': 'lh_free_block_kernel (K1b, source-free)', 'k1s': 'lh_source_block_kernel (K1s)',
                               'k4': 'count_block_kernel (K4)'}[key],
                    'us_per_launch': us, 'us_min': us_min, 'l2': 'flushed before every timed launch',
                    'shape': '64 chains x 2000 x 200 x 8, 7 zones, 20 families'}
    eng.close()
    return out


def run_gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device (the product has no CPU fallback; use --impl reference for the CPU arm)')

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        pool = cpu_pool(cores)                     # forked before CUDA is initialised in this process
        v, dt = cpu_measure(pool, cores, args.cpu_proposals)
        pool.close()
        cpu = {'value': v, 'unit': 'proposals/s', 'cores': cores, 'kind': 'port', 'cpu': cpu_name(),
               'sample': f'{cores} processes x 1 chain x {args.cpu_proposals} proposals on shape D in {dt:.1f} s '
                         f'(oracle port of ZoneMCMC.step)'}
        cpu.update(cpu_likelihood())

    torch.cuda.set_device(local_rank)
    # stdout carries exactly one JSON line: NCCL prints its version banner (NCCL_DEBUG >= VERSION) and other library
    # chatter on fd 1, so everything written to fd 1 from here on goes to stderr and the line is written to the saved fd
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    # torch.cuda.Event only sees torch's current stream: make a real (non-legacy) stream current and hand the
    # same stream to the C ABI, so that events, flushes, NCCL and our kernels are ordered on one stream
    torch.cuda.set_stream(torch.cuda.Stream())
    stream = torch.cuda.current_stream().cuda_stream
    assert stream != 0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rank_max(*vals):
        t = torch.tensor(vals, dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    data = d_data()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    chains_total = args.chains_total
    if chains_total % world:
        raise SystemExit(f'bench.py: {chains_total} chains do not shard evenly over {world} GPUs')

    # ---- device-resident leg: configs[3], 64 MC3 chains in total sharded over the GPUs (strong scaling) ----------
    clocks = ClockSampler(local_rank) if rank == 0 else None
    legs = []
    for seed in args.seeds:
        leg = sampler_leg(torch, args, data, chains_total, seed, stream, barrier, flush, args.steps, args.warmup,
                          keep_states=seed == args.seeds[0])
        leg['step_ms'], leg['kernel_ms'] = rank_max(leg['step_ms'], leg['kernel_ms'])
        legs.append(leg)
    order = sorted(range(len(legs)), key=lambda i: legs[i]['step_ms'])
    med = legs[order[len(order) // 2]]                                     # the median seed
    clock_summary = clocks.summary(min(l['t0'] for l in legs), max(l['t1'] for l in legs)) if clocks else None
    proposals = chains_total * args.mcmc_steps
    value = proposals / (med['step_ms'] * 1e-3)
    acc = np.sum([l['acc'] for l in legs], axis=0)
    rej = np.sum([l['rej'] for l in legs], axis=0)
    tot = torch.tensor(np.stack([acc, rej]).astype(np.float64), device='cuda')
    if world > 1:
        dist.all_reduce(tot)
    acc, rej = tot.cpu().numpy()
    op_names = {0: 'shrink_zone', 1: 'grow_zone', 2: 'swap_zone', 4: 'gibbs_sample_weights', 5: 'gibbs_sample_p_global',
                6: 'gibbs_sample_p_zones', 7: 'gibbs_sample_p_families'}
    n_prop = float((acc + rej).sum())
    acceptance = {k: float(acc[i] / max(1.0, acc[i] + rej[i])) for i, k in op_names.items()}
    sweep_share = float(acc[:3].sum() / max(1.0, n_prop))       # every accepted zone move pays the O(N F) source sweep

    # ---- weak line (N > 1): 64 chains resident PER GPU, the round-1 configuration ----------------------------------
    weak = None
    if world > 1 and not args.no_weak:
        wl = sampler_leg(torch, args, data, chains_total * world, args.seeds[0], stream, barrier, flush, args.steps, args.warmup)
        w_step, w_kernel = rank_max(wl['step_ms'], wl['kernel_ms'])
        weak = {'value': chains_total * world * args.mcmc_steps / (w_step * 1e-3), 'unit': 'proposals/s',
                'chains_per_gpu': chains_total, 'chains_total': chains_total * world, 'ms_per_step': w_step,
                'us_per_proposal_per_chain': 1e3 * w_kernel / args.mcmc_steps, 'scaling': 'weak'}

    # ---- e2e leg: host buffers in, host results out, every step ----------------------------------------------------
    nc_local = chains_total // world
    e2e = e2e_leg(torch, args, data, nc_local, rank * nc_local, local_rank, args.seeds[0], barrier, args.e2e_steps,
                  legs[0].pop('states'))
    e2e_s, e2e_serial_s = rank_max(e2e['s_per_step'], e2e['serial_s_per_step'])
    e2e_value = proposals / e2e_s

    # ---- the drop-in API end to end (sharded ZoneMCMC.generate_samples, 1000 steps per logged sample) --------------
    api = None
    if not args.no_api_leg:
        api = {'shape_d': api_leg(torch, data, chains_total, args.api_steps, args.api_steps // 1000, args.seeds[0])}
        (t_api,) = rank_max(api['shape_d']['sampling_time_s'])
        api['shape_d']['value'] = chains_total * args.api_steps / t_api
        api['shape_d']['vs_device_value'] = api['shape_d']['value'] / value
        if world == 1:
            from contact_zones_b200 import synthetic
            small = synthetic.make_data(28, 47, 3, n_families=5, seed=4, with_dist=False, ragged_states=True,
                                        planted_zones=3, planted_size=5)
            cfg = dict(MODEL_CFG, N_AREAS=3, MAX_M=10)
            api['balkan_size'] = api_leg(torch, small, 4, 10 * args.api_steps, 10 * args.api_steps // 1000, args.seeds[0], cfg)
            api['balkan_size']['shape'] = 'synthetic 28 x 47 x 3, 3 zones, 5 families, 4 MC3 chains'

    # ---- SAMPLE_SOURCE = false: the Metropolis alter_* operators on the same data (SURVEY 8(d) wants both modes) --
    variants = None
    if rank == 0 and not args.no_variants:
        st0 = [d_state(data, 100 + c) for c in range(nc_local)]
        variants = {'sample_source_false': source_free_leg(torch, args, data, st0, local_rank, stream)}

    # ---- roofline legs (rank 0): K1 on configs[4]'s sweep, K1 / K1s / K4 on shape D ---------------------------------
    roof = sweep = kd = None
    if rank == 0 and not args.no_likelihood_leg:
        peak, how = measured_peak()
        sweep = likelihood_leg(torch, local_rank)
        for r in sweep:
            r['frac'] = r['gbs_mean'] / peak
        top = sweep[-1]
        roof = {'bound': 'hbm', 'kernel': 'lh_table_kernel<2> (K1, source-free) on 256 chains x 20000 x 100 x 4',
                'achieved': top['gbs_mean'], 'peak': peak, 'peak_source': how, 'unit': 'GB/s',
                'frac': top['gbs_mean'] / peak, 'traffic': k1_traffic(),
                'bytes_per_launch': top['bytes_per_launch'], 'us_per_launch': top['us_mean'],
                'limiter': 'shared-memory look-up pipe (x is shared by all chains and L2-resident: DRAM traffic is ~1 % of '
                           'the algorithmic bytes, see traffic)',
                'sweep_frac': {str(r['N']): r['frac'] for r in sweep}}
        kd = shape_d_kernel_leg(torch, data, local_rank, stream, peak)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {
        'metric': METRIC, 'value': value, 'unit': 'proposals/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': med['step_ms'], 'higher_is_better': True,
        'scaling': 'strong',          # 64 chains in total whatever N is (the `weak` object holds 64 per GPU)
        'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic', 'config': workload_config(args, world),
        'clocks': clock_summary,
        'e2e': {'value': e2e_value, 'unit': 'proposals/s', 'h2d_bytes_per_step': e2e['h2d'],
                'd2h_bytes_per_step': e2e['d2h'], 'steps': args.e2e_steps, 'upload_ms': e2e['upload_ms'],
                'upload_gbs': e2e['upload_gbs'],
                'serial_value': proposals / e2e_serial_s,
                'pipeline': 'two contexts alternate: the upload of step k+1 overlaps the launch of step k, and the launch '
                            'of step k+1 starts on the SMs that the finished chains of step k free (a launch lasts as long as '
                            'its slowest chain), which is why e2e can exceed the one-stream device rate `value`; '
                            'serial_value = upload -> launch -> read back on one context, nothing overlapped; inputs = '
                            'the warmed chain states of the device leg, re-uploaded from pinned host memory every step'},
        'gpu_launches': int(sum(l['launches'] for l in legs)),
        'kernel_ms_per_step': med['kernel_ms'], 'exchange_ms_per_step': med['exchange_ms'],
        'us_per_proposal_per_chain': 1e3 * med['kernel_ms'] / args.mcmc_steps,
        'chains_per_gpu': nc_local,
        'seeds': [{'seed': l['seed'], 'ms_per_step': l['step_ms'], 'value': proposals / (l['step_ms'] * 1e-3)} for l in legs],
        'll_drift_rel': max(l['drift'] for l in legs),
        'acceptance': acceptance, 'sweep_share_of_steps': sweep_share,
        'mc3_swaps': {'n_swaps': sum(l['swaps']['n_swaps'] for l in legs),
                      'accepted': sum(l['swaps']['accepted_swaps'] for l in legs)},
        'weak': weak, 'api_e2e': api,
        'roofline': roof, 'roofline_k1_shape_d': kd and kd['k1'], 'roofline_k1s': kd and kd['k1s'],
        'roofline_k4': kd and kd['k4'], 'likelihood_sweep': sweep, 'sampler_variants': variants, 'cpu_baseline': cpu,
    }
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + '\n').encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--chains-total', type=int, default=64, help='MC3 chains in total, sharded over the GPUs')
    ap.add_argument('--mcmc-steps', type=int, default=5000, help='proposals per chain per bench step')
    ap.add_argument('--seeds', type=lambda v: [int(x) for x in v.split(',')], default=[0, 1, 2],
                    help='seeds of the device-resident leg (the median is reported)')
    ap.add_argument('--chain-warm-steps', type=int, default=2000, help='untimed steps per chain before the timed region')
    ap.add_argument('--api-steps', type=int, default=100000, help='N_STEPS of the ZoneMCMC.generate_samples leg')
    ap.add_argument('--no-weak', action='store_true')
    ap.add_argument('--no-api-leg', action='store_true')
    ap.add_argument('--e2e-steps', type=int, default=12)
    ap.add_argument('--cpu-proposals', type=int, default=400, help='proposals per host process in the CPU sample')
    ap.add_argument('--seed', type=int, default=1)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-likelihood-leg', action='store_true')
    ap.add_argument('--no-variants', action='store_true')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))
    if args.impl == 'reference':
        run_reference_arm(args, rank)
    else:
        if world != args.gpus and world == 1 and args.gpus > 1:
            raise SystemExit('bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)')
        run_gpu_arm(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
