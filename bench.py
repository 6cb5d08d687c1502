"""bench.py -- the driver's measurement contract for the B200-native sBayes MCMC inner loop.

    python bench.py --gpus N --steps K --warmup W                 (N>1: launched by torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W  (CPU arm: the oracle port on the host cores)

Metric (BASELINE.json): MCMC proposals/s over batched MC3 chains; likelihood GB/s vs the HBM peak.
Workload: configs[3] "synthetic 2000 objects x 200 features x 8 states, 7 zones, 20 families, 64 MC3 chains":
64 chains resident PER GPU (weak scaling: the chains are independent Markov chains, sharded with no data-path
collective; the only exchange is the MC3 swap's all-gather of (ll, prior), 16 B per chain, once per step).

One bench "step" = `--mcmc-steps` proposals on every chain in one persistent-kernel launch (K8) followed by the
MC3 swap exchange (mcmc_generative.py:242-280) -- the unit the reference's `generate_samples` repeats.
  value      proposals/s with every chain state resident in HBM (device-timed, CUDA events, max over ranks)
  e2e        the same through the C-ABI with HOST buffers: every step uploads all chain states from pinned host
             memory (sb_set_state), runs the same launch and reads the logged chain's state + all (ll, prior) back
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
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': workload_config(args, 1),
        'cpu_baseline': {'value': value, 'unit': 'proposals/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'cpu': cpu_name()},
        'e2e': {'value': value, 'unit': 'proposals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0}), flush=True)


def workload_config(args, world):
    return {'workload': 'synthetic 2000x200x8, 7 zones, 20 families, 64 MC3 chains per GPU, SAMPLE_SOURCE, '
                        'default operator mix', 'chains_per_gpu': args.chains, 'chains_total': args.chains * world,
            'mcmc_steps_per_step': args.mcmc_steps, 'swap_exchange': 'all_gather(ll, prior) per step',
            'l2': 'L2 flushed (256 MiB write) between timed steps'}


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
def k1_traffic():
    """DRAM bytes per launch of K1 at N=20000 from the committed `ncu --set full` summary (profiles/), or None."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', 'r*_k1_ncu_full_N20000.txt')))
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
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_launch)]
        for a, b in evs:
            a.record()
            eng.eval_likelihood_async()
            b.record()
        torch.cuda.synchronize()
        ms = np.array([a.elapsed_time(b) for a, b in evs])
        # the same launches back to back inside ONE event pair: the stream never runs dry, so this is pure device time
        # per launch (a per-launch event pair also sees the host's enqueue gap when the kernel is shorter than it)
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
    nc = args.chains
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


def run_gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from contact_zones_b200.engine import Engine
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

    data = d_data()
    nc = args.chains
    eng = Engine(state_idx(data), data.states, data.families, data.network['adj_mat'], None, n_zones=D_SHAPE['Z'],
                 n_chains=nc, inheritance=True, sample_source=True, device=local_rank, chain_offset=rank * nc)
    eng.set_stream(stream)
    states = [d_state(data, 100 + rank * nc + c) for c in range(nc)]
    # pinned host image of all chain states, stacked along the chain axis (the e2e leg uploads from these every step)
    host = {k: torch.from_numpy(np.ascontiguousarray(np.stack(v))).pin_memory() for k, v in dict(
        zones=[s['zones'].astype(np.uint8) for s in states], weights=[s['weights'] for s in states],
        p_global=[s['p_global'][0] for s in states], p_zones=[s['p_zones'] for s in states],
        p_families=[s['p_families'] for s in states],
        source=[np.argmax(s['source'], -1).astype(np.uint8) for s in states]).items()}

    def upload():
        eng.set_states(0, host['zones'].numpy(), host['weights'].numpy(), host['p_global'].numpy(),
                       host['p_zones'].numpy(), host['p_families'].numpy(), host['source'].numpy())
        return sum(t.numel() * t.element_size() for t in host.values())

    upload()
    eng.eval()
    eng.set_operators(default_ops(), PRECISION, min_size=3, max_size=50, p_grow_connected=0.85, seed=args.seed)

    scal = torch.zeros(2 * nc, dtype=torch.float64, device='cuda')
    gathered = torch.zeros(2 * nc * world, dtype=torch.float64, device='cuda')
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')

    def swap_exchange():
        """MC3 swap (mcmc_generative.py:242-280): every rank needs (ll, prior) of every chain."""
        ll, pr = eng.chain_scalars()
        if world > 1:
            scal.copy_(torch.from_numpy(np.concatenate([ll, pr])))
            dist.all_gather_into_tensor(gathered, scal)
            return gathered.cpu().numpy()
        return np.concatenate([ll, pr])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident leg -------------------------------------------------------------------------
    for _ in range(args.warmup):
        eng.run(args.mcmc_steps)
        swap_exchange()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    l0 = eng.kernel_launches
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0 = time.perf_counter()
    for a, b in ev:
        flush.fill_(1)
        a.record()
        eng.run(args.mcmc_steps)
        b.record()
        swap_exchange()
    barrier()
    t1 = time.perf_counter()
    launches = eng.kernel_launches - l0
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    step_ms = ev[0][0].elapsed_time(ev[-1][1]) / args.steps if args.steps > 1 else kernel_ms
    step_ms = max(step_ms, kernel_ms)
    tt = torch.tensor([step_ms, kernel_ms], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    step_ms, kernel_ms = tt.tolist()
    clock_summary = clocks.summary(t0, t1) if clocks else None
    proposals = nc * world * args.mcmc_steps
    value = proposals / (step_ms * 1e-3)

    # drift check of the incrementally maintained ll / prior against a from-scratch evaluation (K1 + K2)
    ll_run, pr_run = eng.chain_scalars()
    ll_new, pr_new, _ = eng.eval()
    drift = float(np.max(np.abs(ll_run - ll_new) / np.abs(ll_new)))
    acc, rej = eng.op_stats()

    # ---- e2e leg: host buffers in, host results out, every step ----------------------------------------
    upload()                    # one untimed pass: the e2e leg starts with warm staging buffers, like the device leg
    eng.run(args.mcmc_steps)
    barrier()
    t0 = time.perf_counter()
    h2d = d2h = 0
    for _ in range(args.e2e_steps):
        h2d = upload()
        eng.run(args.mcmc_steps)
        s0 = eng.get_state(0, one_hot_source=False)
        ll, pr = eng.chain_scalars()
        d2h = sum(v.nbytes for v in s0.values() if v is not None) + ll.nbytes + pr.nbytes
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.e2e_steps
    te = torch.tensor([e2e_s], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = proposals / te.item()
    eng.close()

    # ---- SAMPLE_SOURCE = false: the Metropolis alter_* operators on the same data (SURVEY 8(d) wants both modes) --
    variants = None
    if rank == 0 and not args.no_variants:
        variants = {'sample_source_false': source_free_leg(torch, args, data, states, local_rank, stream)}

    # ---- likelihood roofline leg (rank 0) ---------------------------------------------------------------
    roof = sweep = None
    if rank == 0 and not args.no_likelihood_leg:
        sweep = likelihood_leg(torch, local_rank)
        top = sweep[-1]
        peak, how = measured_peak()
        roof = {'bound': 'hbm', 'kernel': 'lh_table_kernel<2> (K1, source-free) on 256 chains x 20000 x 100 x 4',
                'achieved': top['gbs_mean'], 'peak': peak, 'peak_source': how, 'unit': 'GB/s',
                'frac': top['gbs_mean'] / peak, 'traffic': k1_traffic(),
                'bytes_per_launch': top['bytes_per_launch'], 'us_per_launch': top['us_mean']}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {
        'metric': METRIC, 'value': value, 'unit': 'proposals/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': step_ms, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic', 'config': workload_config(args, world),
        'clocks': clock_summary,
        'e2e': {'value': e2e_value, 'unit': 'proposals/s', 'h2d_bytes_per_step': int(h2d) * 1,
                'd2h_bytes_per_step': int(d2h), 'steps': args.e2e_steps},
        'gpu_launches': int(launches),
        'kernel_ms_per_step': kernel_ms, 'us_per_proposal_per_chain': 1e3 * kernel_ms / args.mcmc_steps,
        'll_drift_rel': drift,
        'acceptance': {k: float(acc[:, i].sum() / max(1, acc[:, i].sum() + rej[:, i].sum()))
                       for k, i in (('shrink_zone', 0), ('grow_zone', 1), ('swap_zone', 2))},
        'roofline': roof, 'likelihood_sweep': sweep, 'sampler_variants': variants, 'cpu_baseline': cpu,
    }
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + '\n').encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--chains', type=int, default=64, help='MC3 chains resident per GPU')
    ap.add_argument('--mcmc-steps', type=int, default=2000, help='proposals per chain per bench step')
    ap.add_argument('--e2e-steps', type=int, default=5)
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
