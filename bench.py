#!/usr/bin/env python
"""bench.py -- model-rollout transitions/s of the 7-net ensemble (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload halfcheetah|hopper|walker2d] [--precision fp32|bf16] [--no-extras]

A "step" is ONE whole model rollout: B=50 000 starts x h=5 fused FakeEnv steps (ensemble forward,
sampling, elite gather, penalty, termination, compaction) appended to a device-resident 1.25 M-row
model_pool -- the HalfCheetah-shaped configuration north_star sets its target on (BASELINE.json
configs[0] shape / SURVEY.md 8(d) config 1).  Prints ONE JSON line.

The headline arm runs at the reference's arithmetic: MOPO_PREC_FP32, the fp32-accurate tcgen05 kernel (every operand an
fp16 (hi, lo) pair, three MMAs per K step, fp32 accumulation; passes the rtol-1e-3 parity tests).  The plain-bf16
tensor-core path is reported beside it (`bf16_path`) together with its MEASURED error against the float64 oracle.

Timing: W >= 3 warm-up rollouts, then K rollouts, each bracketed by CUDA events on the launching
stream with a 256 MiB L2 flush between rollouts (outside the event pairs); barrier + synchronize
on both sides; max over ranks.  `value` has inputs resident in HBM; `e2e` goes through the public
Python API with pinned HOST buffers (H2D of start states and actions, D2H of the read-back, inside
the timed region).  `roofline` is the ensemble-forward kernel, CUDA events around every launch inside K
more rollouts, against the measured bf16 tensor peak (burst when the run was a short un-capped burst,
else sustained; / 3 for the three-MMA fp32 path).  `cpu_baseline` is the oracle (NumPy port of the
reference CPU path; TensorFlow is not installed) at the SAME configuration on this box's host cores.
Extra records in the same line: `bf16_path`, `path_error`, `configs` (BASELINE.json configs 2, 3, 5) and
`strong_scaling` (config 4: 1 M starts sharded over the N GPUs, with the NCCL parameter broadcast and the
pool all-gather timed).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {'halfcheetah': (17, 6), 'hopper': (11, 3), 'walker2d': (17, 6)}
E, H_DIM, N_HIDDEN, ELITES = 7, 200, 4, 5
METRIC = 'model-rollout transitions/sec (7-net ensemble)'
UNIT = 'transitions/s'
DTYPE_NOTE = {'fp32': 'fp32-accurate: fp16 (hi, lo) operand pairs, 3 tcgen05 MMAs per K step, fp32 accumulate in TMEM (MOPO_PREC_FP32)',
              'bf16': 'bf16 operands, fp32 accumulate in TMEM (MOPO_PREC_BF16)'}


def flop_per_transition(obs_dim, act_dim):
    din, d1 = obs_dim + act_dim, obs_dim + 1
    return 2 * E * (din * H_DIM + (N_HIDDEN - 1) * H_DIM * H_DIM + 2 * H_DIM * d1)


def synthetic(workload, B, h, seed=0):
    """Reference init law weights (fc.py:145-154), N(0,1)-shaped starts, tanh(N(0,1)) actions, fitted scaler."""
    from oracle import ensemble as ens
    obs_dim, act_dim = WORKLOADS[workload]
    p = ens.init_params(obs_dim, act_dim, seed=seed, separate_mean_var=True)
    rng = np.random.RandomState(seed + 1)
    obs = rng.standard_normal((B, obs_dim)).astype(np.float32)
    if workload in ('hopper', 'walker2d'):
        obs *= 0.05
        obs[:, 0] += 1.25
        p['max_logvar'][:] = -6.0            # "low-noise" variant so most rows survive 5 steps
    acts = [np.tanh(rng.standard_normal((B, act_dim))).astype(np.float32) for _ in range(h)]
    ens.fit_scaler(p, np.concatenate([obs, acts[0]], -1))
    return p, obs, acts


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = 'index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,' \
        'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'

    def __init__(self, gpu):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith('active'):
                        reasons.add(n)
            except Exception:
                pass
        busy = sorted(s for s in sm if s > 0)
        return {'sm_mhz': busy[len(busy) // 2] if busy else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        d = json.load(open(path))
        return d.get('bf16_tflops', 1590.0), d.get('bf16_tflops_sustained', 1400.0), d.get('hbm_gbs', 6650.0), 'measured'
    return 1590.0, 1400.0, 6650.0, 'fallback'


# --------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the oracle = NumPy port of the reference CPU path
# --------------------------------------------------------------------------------------------------
def cpu_rollout_rate(workload, B_sample, h, reps, warmup=1, budget_s=None):
    """The literal _rollout_model loop of the reference on the host cores (oracle port: BLAS-threaded NumPy fp32 forward,
    restated FakeEnv.step / ring pool), BLAS pinned to every core.  budget_s: stop adding timed repetitions once the
    budget is spent (at least one is always timed).  Returns (transitions/s, median seconds, transitions, reps timed)."""
    from oracle import ensemble as ens
    from oracle.fake_env import FakeEnv
    from oracle.static_fns import StaticFns
    from oracle.pool import SimpleReplayPool
    from oracle.rollout import rollout_model
    try:
        from threadpoolctl import threadpool_limits
        limiter = threadpool_limits(limits=os.cpu_count())
    except Exception:
        limiter = None
    obs_dim, act_dim = WORKLOADS[workload]
    p, obs, acts = synthetic(workload, B_sample, h)
    env = FakeEnv(ens.OracleModel(p), StaticFns(workload), penalty_coeff=1.0, penalty_learned_var=True)
    pool = SimpleReplayPool(obs_dim, act_dim, 1250000)
    np.random.seed(0)
    times, trans = [], 0
    t_start = time.perf_counter()
    for i in range(warmup + reps):
        t0 = time.perf_counter()
        _, steps = rollout_model(env, pool, obs, h, lambda k, o: acts[k][:len(o)])
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt); trans = sum(steps)
            if budget_s is not None and time.perf_counter() - t_start > budget_s:
                break
    if limiter is not None:
        limiter.unregister() if hasattr(limiter, 'unregister') else None
    med = sorted(times)[len(times) // 2]
    return trans / med, med, trans, len(times)


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.
    TensorFlow 1.14 is not installed, so this is the oracle port (kind 'port'): NumPy fp32 forward with
    BLAS-threaded matmul + the restated FakeEnv.step / ring pool / _rollout_model loop, at the SAME configuration as
    our arm (full rollout_batch_size); the number of timed repetitions is bounded by a wall-clock budget."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count()
    B_sample = args.cpu_sample if args.cpu_sample > 0 else args.batch
    rate, med, trans, reps = cpu_rollout_rate(args.workload, B_sample, args.h, reps=max(args.steps, 1), warmup=1, budget_s=150.0)
    obs_dim, act_dim = WORKLOADS[args.workload]
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': med * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'fp32', 'data': 'synthetic',
        'config': {'workload': '%s-shaped (obs %d, act %d) 7-net/5-elite 4x200 ensemble rollout, B=%d starts, h=%d, lambda=1.0' % (args.workload, obs_dim, act_dim, args.batch, args.h),
                   'rollout_batch_size': args.batch, 'rollout_length': args.h},
        'cpu_baseline': {'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'blas_threads': cores,
                         'same_config': B_sample == args.batch,
                         'sample': 'oracle NumPy rollout, B=%d starts x h=%d (%d transitions per timed step), median of %d timed rollouts (150 s budget)' % (B_sample, args.h, trans, reps)},
        'e2e': {'value': rate, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide handles of one bench run."""
    pass


def quiet(fn, *a, **k):
    """The shims print like the reference; keep stdout to ONE json line."""
    import contextlib
    with open(os.devnull, 'w') as devnull, contextlib.redirect_stdout(devnull):
        return fn(*a, **k)


def make_case(cx, workload, precision, B, h, pool_rows=None, policy=False, seed_rank=True):
    """Model + FakeEnv + ModelRollout + resident and pinned-host inputs for one (workload, precision, B, h)."""
    import torch
    from mopo_b200 import _lib, dist as mdist
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static
    obs_dim, act_dim = WORKLOADS[workload]
    prec = {'bf16': _lib.PREC_BF16, 'fp32': _lib.PREC_FP32}[precision]
    c = Ctx()
    c.workload, c.precision, c.prec, c.B, c.h, c.obs_dim, c.act_dim = workload, precision, prec, B, h, obs_dim, act_dim
    p, obs, acts = synthetic(workload, B, h, seed=cx.rank if seed_rank else 0)        # per-rank starts, same weights below
    p0, _, _ = synthetic(workload, 8, h, seed=0)
    p0['mu'], p0['sigma'] = p['mu'], p['sigma']
    c.params = p0
    model = construct_model(obs_dim=obs_dim, act_dim=act_dim, hidden_dim=H_DIM, num_networks=E, num_elites=ELITES,
                            separate_mean_var=True, name='bench', device=cx.local, precision=prec)
    model.set_params(p0['weights'], p0['biases'], p0['var_weights'], p0['var_biases'], p0['mu'], p0['sigma'],
                     p0['max_logvar'], p0['min_logvar'])
    model._model_inds = list(range(ELITES))
    c.model = model
    c.env = FakeEnv(model, static[workload], penalty_coeff=1.0, penalty_learned_var=True, seed=123)
    env_pool = SimpleReplayPool(obs_dim, act_dim, 16, device=cx.local)
    c.pol = None
    if policy:                                     # act = policy(obs) on the device every step
        from mopo_b200.policies import GaussianPolicy
        from oracle import policy as opol          # (random-init weights only; nothing of oracle/ is executed in the timed path)
        c.pol = GaussianPolicy(obs_dim, act_dim, device=cx.local, precision=prec, seed=123)
        c.pol.set_params(*opol.init_params(obs_dim, act_dim, seed=0))
    retain = 5 if pool_rows is None else max(1, pool_rows // (B * h))
    c.ro = ModelRollout(c.env, env_pool, rollout_batch_size=B, rollout_schedule=(20, 100, h, h), epoch_length=1000,
                        model_train_freq=1000, model_retain_epochs=retain, seed=123, policy=c.pol)
    c.ro._set_rollout_length()
    c.ro._reallocate_model_pool()                  # 5 * h * B = 1.25 M rows at the default config (d4rl/base.py:19-21)
    c.d_obs = torch.as_tensor(obs, device=cx.dev)
    c.d_acts = [torch.as_tensor(a, device=cx.dev) for a in acts]
    c.h_obs = torch.as_tensor(obs).pin_memory()
    c.h_acts = [torch.as_tensor(a).pin_memory() for a in acts]
    c.h_rew = torch.empty(B, 1, dtype=torch.float32).pin_memory()
    c.fpt = flop_per_transition(obs_dim, act_dim)
    c.origin = cx.rank * B
    return c


def timed(cx, fn, steps, warmup, sample_clocks=False):
    import torch
    from mopo_b200 import dist as mdist
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(cx.dev)
    mdist.barrier()
    clk = ClockSampler(cx.local) if sample_clocks else None
    if clk:
        clk.start()
    evs = []
    trans = 0
    l0 = cx.lib.mopo_launch_count()
    for _ in range(steps):
        cx.flush.fill_(1)                                                  # L2 flush, outside the event pair
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(cx.stream)
        trans += fn()
        e.record(cx.stream)
        evs.append((s, e))
    torch.cuda.synchronize(cx.dev)
    l1 = cx.lib.mopo_launch_count()
    clocks = clk.stop() if clk else None
    mdist.barrier()
    ms = sum(s.elapsed_time(e) for s, e in evs)
    return ms, trans, l1 - l0, clocks


def whole_job(cx, ms, trans):
    """transitions of all ranks / max-over-ranks time."""
    import torch
    from mopo_b200 import dist as mdist
    ms_max = mdist.max_over_ranks(ms, device=cx.dev)
    tot = torch.tensor([float(trans)], dtype=torch.float64, device=cx.dev)
    if cx.world > 1:
        torch.distributed.all_reduce(tot)
    return float(tot.item()) / (ms_max * 1e-3), ms_max


def measure(cx, c, steps, warmup, clocks=False, kernel=True, e2e=True):
    """Resident value, e2e and the forward kernel's roofline of one case.  Every rank runs it; the numbers are whole-job."""
    import torch
    from mopo_b200 import _lib
    lib, ro, B, h = cx.lib, c.ro, c.B, c.h

    def rollout_resident():
        ro._rollout_model(B, start_obs=c.d_obs, actions=None if c.pol else c.d_acts, origin_base=c.origin, quiet=True)
        return sum(ro._last_steps_added)

    out = {}
    for _ in range(warmup):
        rollout_resident()
    # pass 1: the headline `value` (K rollouts, CUDA events around each whole rollout, nothing recorded inside)
    ms, trans, launches, clk = timed(cx, rollout_resident, steps, 0, sample_clocks=clocks)
    out['value'], ms_max = whole_job(cx, ms, trans)
    out.update(ms_per_step=ms_max / steps, launches=launches, clocks=clk, trans_per_step=trans // steps, timed_ms=ms_max)
    if kernel:
        # passes 2, 3: the same K rollouts again with CUDA events recorded on the launching stream around every launch of
        # the forward kernel (-> roofline) and of the step tail (-> what the rest of the step costs)
        kt_ms, kt_n, t_ms, t_n = C.c_double(0), C.c_int64(0), C.c_double(0), C.c_int64(0)
        _lib.check(lib.mopo_kernel_timing(c.model._handle, 1))
        timed(cx, rollout_resident, steps, 0)
        _lib.check(lib.mopo_kernel_time(c.model._handle, C.byref(kt_ms), C.byref(kt_n)))
        _lib.check(lib.mopo_kernel_timing(c.model._handle, 2))
        timed(cx, rollout_resident, steps, 0)
        _lib.check(lib.mopo_kernel_time(c.model._handle, C.byref(t_ms), C.byref(t_n)))
        _lib.check(lib.mopo_kernel_timing(c.model._handle, 0))
        if kt_n.value > 0:
            ms_f = kt_ms.value / kt_n.value
            rows = trans / (h * steps)
            burst, sustained, hbm, src = peaks()
            capped = bool(clk and 'sw_power_cap' in (clk.get('reasons') or []))
            use_burst = (not capped) and ms_max < 1000.0
            mma_per_mac = 3 if c.precision == 'fp32' else 1
            peak = (burst if use_burst else sustained) / mma_per_mac
            achieved = c.fpt * rows / (ms_f * 1e-3) / 1e12
            kname = 'ens_tc3_kernel<3>' if c.precision == 'fp32' else 'ens_tc_kernel'
            traffic = None
            tpath = os.path.join(ROOT, 'profiles', 'roofline_traffic.json')
            if os.path.exists(tpath):
                traffic = json.load(open(tpath)).get(kname, {}).get('dram_bytes_per_launch')
            out['roofline'] = {
                'bound': 'tensor', 'achieved': achieved, 'peak': peak, 'unit': 'TFLOP/s', 'frac': achieved / peak, 'traffic': traffic,
                'kernel': '%s (tcgen05 ensemble forward + fused head, %s), %d rows/launch; timed inside %d rollouts of the same workload (%d launches, CUDA events on the launching stream)'
                          % (kname, c.precision, rows, steps, kt_n.value),
                'kernel_ms': ms_f, 'flop_per_transition': c.fpt,
                'peak_source': '%s bf16 %s peak %.1f TFLOP/s%s (timed region %.0f ms, sw_power_cap %s)' % (
                    src, 'burst' if use_burst else 'sustained', burst if use_burst else sustained,
                    ' / 3: the fp32-accurate path issues three kind::f16 MMAs per algorithmic multiply-add' if mma_per_mac == 3 else '',
                    ms_max, 'seen' if capped else 'not seen'),
                'frac_of_burst': achieved / (burst / mma_per_mac), 'frac_of_sustained': achieved / (sustained / mma_per_mac),
                'tensor_pipe_tflops_issued': achieved * mma_per_mac,
                'kernel_share_of_step': (ms_f * h * steps) / ms, 'tail_kernel_ms': t_ms.value / max(t_n.value, 1),
                'rollout_frac_of_peak': out['value'] / cx.world * c.fpt / 1e12 / peak}
    if e2e:
        h2d = c.h_obs.numel() * 4 + (0 if c.pol else sum(a.numel() * 4 for a in c.h_acts))
        d2h = c.h_rew.numel() * 4

        import time as _t
        prof = []

        def rollout_e2e():
            _a = _t.perf_counter()
            # pinned host buffers straight into the public call: it uploads the starts + step 0's actions on the compute
            # stream and the later steps' actions on a side stream, overlapped with the earlier steps
            ro._rollout_model(B, start_obs=c.h_obs, actions=None if c.pol else c.h_acts, origin_base=c.origin, quiet=True)
            _b = _t.perf_counter()
            pool = ro._model_pool                           # read-back: the penalised rewards of the rollout's last step
            n_last = ro._last_steps_added[-1] if ro._last_steps_added else 0
            last = (pool._pointer - n_last) % pool._max_size
            if n_last and last + n_last <= pool._max_size:
                c.h_rew[:n_last].copy_(pool.fields['rewards'][last:last + n_last], non_blocking=True)
            elif n_last:
                k = pool._max_size - last
                c.h_rew[:k].copy_(pool.fields['rewards'][last:], non_blocking=True)
                c.h_rew[k:n_last].copy_(pool.fields['rewards'][:n_last - k], non_blocking=True)
            _c = _t.perf_counter()
            torch.cuda.current_stream(cx.dev).synchronize()
            _d = _t.perf_counter()
            prof.append((_b - _a, _c - _b, _d - _c))
            return sum(ro._last_steps_added)

        ms_e, trans_e, _, _ = timed(cx, rollout_e2e, steps, warmup)
        if os.environ.get('MOPO_BENCH_PROF'):
            import numpy as _np
            sys.stderr.write('e2e cpu profile (ms): enqueue %.3f  readback-enqueue %.3f  sync-wait %.3f\n' % tuple(_np.median(_np.array(prof[-steps:]), 0) * 1e3))
        v, ms_e_max = whole_job(cx, ms_e, trans_e)
        out['e2e'] = {'value': v, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h, 'ms_per_step': ms_e_max / steps,
                      'result': 'the rollout result (the pool rows) stays in HBM for the SAC sampler, as the reference returns only '
                                "{'mean_rollout_length'} (mopo.py:760-761); per-step counts + the last step's penalised rewards are read back"}
    return out


def path_error(cx, c, n=4096):
    """Measured error of one FakeEnv.step of this case's device path against the float64 oracle, injected noise / elites."""
    from oracle import ensemble as ens
    from oracle.fake_env import FakeEnv as OracleEnv
    from oracle.static_fns import StaticFns
    rng = np.random.RandomState(7)
    obs = c.d_obs[:n].cpu().numpy()
    act = c.d_acts[0][:n].cpu().numpy()
    n = len(obs)
    noise = rng.standard_normal((E, n, c.obs_dim + 1))
    inds = rng.choice(ELITES, size=n)
    o_env = OracleEnv(ens.OracleModel(c.params, dtype=np.float64), StaticFns(c.workload), penalty_coeff=1.0, penalty_learned_var=True)
    r_nobs, r_rew, r_term, r_info = o_env.step(obs.astype(np.float64), act.astype(np.float64), noise=noise, model_inds=inds)
    nobs, rew, term, info = quiet(c.env.step, obs, act, noise=noise, model_inds=inds, precision=c.prec, diagnostics=False)

    def err(a, b):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        d = np.abs(a - b)
        return {'max_abs': float(d.max()), 'max_rel_to_scale': float(d.max() / max(np.abs(b).max(), 1e-30)),
                'outside_rtol1e-3_atol2e-5': int((d > 2e-5 + 1e-3 * np.abs(b)).sum()), 'elements': int(d.size)}
    return {'rows': n, 'vs': 'float64 oracle of bnn.py:631-675 + fake_env.py:37-131, same noise and elite picks',
            'next_obs': err(nobs, r_nobs), 'reward': err(rew, r_rew), 'penalty': err(info['penalty'], r_info['penalty']),
            'elite_mismatches': int((np.asarray(info['model_inds']) != inds).sum()),
            'termination_flips': int((np.asarray(term).ravel() != np.asarray(r_term).ravel()).sum())}


def brief(r):
    """Sub-record: value, e2e, the kernel's roofline fraction."""
    out = {'value': r['value'], 'unit': UNIT, 'ms_per_step': r['ms_per_step'], 'transitions_per_step_per_gpu': r['trans_per_step']}
    if 'e2e' in r:
        out['e2e'] = r['e2e']['value']
    if 'roofline' in r:
        out['roofline'] = {k: r['roofline'][k] for k in ('frac', 'achieved', 'peak', 'kernel_ms', 'peak_source')}
    return out


def epoch_loop(cx, B, h, n_upd, epochs):
    """BASELINE.json config 5: per epoch one rollout (B starts per GPU x h, device actor) into the model pool, the NCCL
    all-gather of every rank's new pool rows (N > 1), n_upd SAC updates (batch 256 = 12 env + 244 model rows drawn on the
    device, replicated with equal seeds) and the actor export back into the rollout policy."""
    import torch
    from mopo_b200 import _lib, dist as mdist
    import mopo_b200.static as static
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    from mopo_b200.algorithms.sac import SACUpdate
    from mopo_b200.policies import GaussianPolicy
    od, ad = WORKLOADS['halfcheetah']
    p, obs, acts = synthetic('halfcheetah', 100000, 1, seed=0)
    rng = np.random.RandomState(1)
    env_pool = SimpleReplayPool(od, ad, 100000, device=cx.local)
    env_pool.add_samples({'observations': obs, 'actions': acts[0], 'next_observations': obs + 0.1 * rng.standard_normal(obs.shape).astype(np.float32),
                          'rewards': rng.standard_normal((100000, 1)).astype(np.float32), 'terminals': np.zeros((100000, 1), bool)})
    model = construct_model(obs_dim=od, act_dim=ad, separate_mean_var=True, device=cx.local, precision=_lib.PREC_FP32)
    model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    model._model_inds = list(range(ELITES))
    env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
    pol = GaussianPolicy(od, ad, device=cx.local, precision=_lib.PREC_FP32, seed=5)
    sac = SACUpdate(od, ad, device=cx.local, seed=7)
    sac.export_policy(pol)
    ro = ModelRollout(env, env_pool, rollout_batch_size=B, rollout_schedule=(0, 1, h, h), epoch_length=1000, model_train_freq=1000,
                      model_retain_epochs=5 if cx.world == 1 else 1, policy=pol, seed=9)
    ro._set_rollout_length()
    ro._reallocate_model_pool()
    shard_pool = ro._model_pool
    full_pool = shard_pool if cx.world == 1 else SimpleReplayPool(od, ad, 5 * B * h * cx.world, device=cx.local)

    def epoch():
        ro._rollout_model(B, quiet=True, origin_base=cx.rank * B)
        if cx.world > 1:
            mdist.allgather_new_rows(shard_pool, full_pool, B * h)
        sac.train_from_pools(env_pool, full_pool, n_updates=n_upd)
        sac.export_policy(pol)

    epoch()
    torch.cuda.synchronize(cx.dev)
    mdist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    l0 = cx.lib.mopo_launch_count()
    ev[0].record(cx.stream)
    for _ in range(epochs):
        epoch()
    ev[1].record(cx.stream)
    torch.cuda.synchronize(cx.dev)
    launches = (cx.lib.mopo_launch_count() - l0) // epochs
    ev[2].record(cx.stream)
    sac.train_from_pools(env_pool, full_pool, n_updates=n_upd)
    ev[3].record(cx.stream)
    torch.cuda.synchronize(cx.dev)
    mdist.barrier()
    ms_epoch = mdist.max_over_ranks(ev[0].elapsed_time(ev[1]) / epochs, device=cx.dev)
    return {'workload': 'MOPO epoch loop: rollout B=%d starts per GPU x h=%d with the device actor (fp32-accurate path) + %d SAC updates (batch 256 = 12 env + 244 model rows) + actor export%s'
                        % (B, h, n_upd, '' if cx.world == 1 else ' + NCCL all-gather of the new pool rows'),
            'ms_per_epoch': ms_epoch, 'ms_per_sac_update': ev[2].elapsed_time(ev[3]) / n_upd,
            'rollout_transitions_per_epoch': B * h * cx.world, 'gpu_launches_per_epoch': launches, 'n_gpus': cx.world}


def strong_scaling(cx, precision, total, h, steps):
    """BASELINE.json config 4: `total` starts sharded over the N GPUs (contiguous ranges, global Philox row ids), with the
    collectives north_star names timed: the NCCL broadcast of the ensemble parameters (as after model training) and the
    all-gather of every rank's new model-pool rows into each rank's pool copy (the SAC sampler's view)."""
    import torch
    from mopo_b200 import dist as mdist
    lo, hi = mdist.shard_range(total, cx.rank, cx.world)
    Bl = hi - lo
    c = quiet(make_case, cx, 'halfcheetah', precision, Bl, h, pool_rows=Bl * h)
    c.origin = lo
    r = measure(cx, c, steps, 3, clocks=False, kernel=True, e2e=True)
    out = {'workload': 'HalfCheetah-shaped, %d starts x h=%d sharded over %d GPU(s) (%d per GPU), %s path' % (total, h, cx.world, Bl, precision),
           'value': r['value'], 'e2e': r['e2e']['value'], 'unit': UNIT, 'ms_per_rollout': r['ms_per_step'],
           'roofline_frac': r.get('roofline', {}).get('frac'), 'kernel_ms': r.get('roofline', {}).get('kernel_ms')}
    if cx.world > 1:
        from mopo_b200.replay_pools import SimpleReplayPool
        # (1) parameter broadcast, device to device
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        mdist.broadcast_params(c.model, src=0)
        torch.cuda.synchronize(cx.dev); mdist.barrier()
        ev[0].record(cx.stream)
        for _ in range(5):
            mdist.broadcast_params(c.model, src=0, sync_host=False)     # the device path alone: pack, NCCL broadcast, unpack, operand images rebuilt
        ev[1].record(cx.stream)
        torch.cuda.synchronize(cx.dev)
        out['broadcast_ms'] = mdist.max_over_ranks(ev[0].elapsed_time(ev[1]) / 5, device=cx.dev)
        out['broadcast_bytes'] = int(c.model.param_bytes())
        out['broadcast_note'] = 'fp32 master images packed device to device, one NCCL broadcast, unpacked in place, tcgen05 operand images rebuilt on the receivers; the host mirrors used by save() are refreshed outside this timing'
        # (2) pool all-gather of one rollout's rows
        full = SimpleReplayPool(c.obs_dim, c.act_dim, total * h, device=cx.local)
        row_bytes = 4 * (2 * c.obs_dim + c.act_dim + 1) + 1
        mdist.allgather_new_rows(c.ro._model_pool, full, Bl * h)
        torch.cuda.synchronize(cx.dev); mdist.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record(cx.stream)
        for _ in range(3):
            mdist.allgather_new_rows(c.ro._model_pool, full, Bl * h)
        ev[1].record(cx.stream)
        torch.cuda.synchronize(cx.dev)
        ag = mdist.max_over_ranks(ev[0].elapsed_time(ev[1]) / 3, device=cx.dev)
        out['pool_allgather_ms'] = ag
        out['pool_allgather_bytes_received_per_gpu'] = int((total - Bl) * h * row_bytes)
        out['pool_allgather_gbs_per_gpu'] = (total - Bl) * h * row_bytes / (ag * 1e-3) / 1e9
        out['nvlink_peer_copy_reference_gbs'] = 770.0
        trans = total * h
        out['e2e_with_collectives'] = trans / (trans / r['e2e']['value'] + (out['broadcast_ms'] + ag) * 1e-3)
    return out


def run_ours(args):
    if os.environ.get('NCCL_DEBUG', 'VERSION').upper() == 'VERSION':
        os.environ['NCCL_DEBUG'] = 'WARN'
    # stdout carries exactly ONE json line: NCCL prints its version banner there from C (first communicator), the shims print
    # like the reference -- so file descriptor 1 points at /dev/null until the line is ready
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    devnull_fd = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull_fd, 1)
    import torch
    from mopo_b200 import _lib, dist as mdist
    import __graft_entry__ as ge
    cx = Ctx()
    cx.rank, cx.world, cx.local = mdist.init_from_env()
    if cx.rank == 0:
        ge.build()
    mdist.barrier()
    torch.cuda.set_device(cx.local)
    cx.dev = torch.device('cuda', cx.local)
    cx.lib = _lib.load()
    cx.flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)
    cx.stream = torch.cuda.current_stream(cx.dev)
    B, h = args.batch, args.h
    obs_dim, act_dim = WORKLOADS[args.workload]

    try:
        main_case = quiet(make_case, cx, args.workload, args.precision, B, h, policy=args.policy == 'device')
        if cx.world > 1:
            mdist.broadcast_params(main_case.model, src=0)                     # NCCL: replicate the ensemble (not timed here)
        quiet(main_case.ro._rollout_model, B, start_obs=main_case.d_obs, actions=None if main_case.pol else main_case.d_acts,
              origin_base=main_case.origin, quiet=True)
    except NotImplementedError as e:
        if cx.rank == 0:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            print(json.dumps({'error': 'precision %s unavailable: %s' % (args.precision, e)}))
        return
    r = measure(cx, main_case, args.steps, args.warmup, clocks=True)
    extras = {}
    if not args.no_extras:
        # every rank runs the same sequence (the records are whole-job numbers); a failing extra is reported in the line, it
        # does not take the headline record with it
        def extra(name, fn):
            try:
                return fn()
            except Exception as e:                                  # noqa: BLE001
                import traceback
                sys.stderr.write('bench extra %s failed:\n%s' % (name, traceback.format_exc()))
                return {'error': '%s: %s' % (type(e).__name__, str(e)[:200])}

        ksteps = max(5, min(args.steps, 10))
        other = 'bf16' if args.precision == 'fp32' else 'fp32'
        extras['path_error'] = {args.precision: extra('path_error', lambda: path_error(cx, main_case))}

        def other_path():
            oc = quiet(make_case, cx, args.workload, other, B, h, policy=args.policy == 'device')
            rec = dict(brief(measure(cx, oc, ksteps, 3)), dtype=DTYPE_NOTE[other])
            extras['path_error'][other] = path_error(cx, oc)
            return rec
        extras['%s_path' % other] = extra('%s_path' % other, other_path)
        cfgs = {}
        for name, wl, prec, hh in (('hopper_h5', 'hopper', 'fp32', 5), ('hopper_h5_bf16', 'hopper', 'bf16', 5),
                                   ('walker2d_h1_parity', 'walker2d', 'fp32', 1), ('walker2d_h1_bf16', 'walker2d', 'bf16', 1)):
            def one(wl=wl, prec=prec, hh=hh):
                cc = quiet(make_case, cx, wl, prec, B, hh)
                rec = brief(measure(cx, cc, ksteps, 3))
                rec['path_error'] = path_error(cx, cc)
                return rec
            cfgs[name] = extra(name, one)
        cfgs['epoch_loop'] = extra('epoch_loop', lambda: quiet(epoch_loop, cx, B, 5, 1000, 2))
        extras['configs'] = cfgs
        extras['strong_scaling'] = extra('strong_scaling', lambda: strong_scaling(cx, args.precision, 1000000, 5, 5))
    if cx.rank != 0:
        return
    cores = os.cpu_count()
    cpu = None
    if not args.no_cpu_baseline:
        Bs = args.cpu_sample if args.cpu_sample > 0 else B
        rate, med, tr, reps = cpu_rollout_rate(args.workload, Bs, h, reps=3, budget_s=40.0)
        cpu = {'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'blas_threads': cores, 'same_config': Bs == B,
               'sample': 'oracle NumPy rollout (BLAS-threaded fp32 forward + restated FakeEnv.step/pool), B=%d x h=%d = %d transitions per rollout, median of %d (TensorFlow absent: reference arm = port)' % (Bs, h, tr, reps)}
    line = {
        'metric': METRIC, 'value': r['value'], 'unit': UNIT, 'n_gpus': cx.world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': r['ms_per_step'], 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': args.precision, 'dtype_note': DTYPE_NOTE[args.precision], 'data': 'synthetic',
        'config': {'workload': '%s-shaped (obs %d, act %d) 7-net/5-elite 4x200 ensemble rollout, B=%d starts per GPU, h=%d, lambda=1.0, into a %d-row device model_pool'
                               % (args.workload, obs_dim, act_dim, B, h, main_case.ro._model_pool._max_size),
                   'rollout_batch_size': B, 'rollout_length': h, 'transitions_per_step_per_gpu': r['trans_per_step'],
                   'actions': 'on-device policy (2x256 ReLU Gaussian actor) every step' if main_case.pol else 'precomputed tanh(N(0,1)) arrays',
                   'l2': 'flushed (256 MiB write) between timed rollouts', 'parallelism': 'batch-sharded x%d, no data-path collective' % cx.world},
        'e2e': r['e2e'],
        'gpu_launches': r['launches'],
        'clocks': r['clocks'],
        'roofline': r.get('roofline'),
        'cpu_baseline': cpu,
    }
    line.update(extras)
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line))
    sys.stdout.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='halfcheetah', choices=sorted(WORKLOADS))
    ap.add_argument('--precision', default=os.environ.get('MOPO_BENCH_PRECISION', 'fp32'), choices=['bf16', 'fp32'])
    ap.add_argument('--batch', type=int, default=50000)
    ap.add_argument('--h', type=int, default=5)
    ap.add_argument('--cpu-sample', type=int, default=0, help='starts in the CPU-baseline rollout (0 = the full --batch: same configuration)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true', help='only the headline record (no bf16_path / path_error / configs / strong_scaling)')
    ap.add_argument('--policy', default='none', choices=['none', 'device'], help="'device': actions from the on-device SAC actor each step (row f1) instead of precomputed action arrays")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)
        from mopo_b200 import dist as mdist
        mdist.shutdown()


if __name__ == '__main__':
    main()
