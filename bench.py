#!/usr/bin/env python
"""bench.py -- model-rollout transitions/s of the 7-net ensemble (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload halfcheetah|hopper|walker2d] [--precision bf16|fp32]

A "step" is ONE whole model rollout: B=50 000 starts x h=5 fused FakeEnv steps (ensemble forward,
sampling, elite gather, penalty, termination, compaction) appended to a device-resident 1.25 M-row
model_pool -- the HalfCheetah-shaped configuration north_star sets its target on (BASELINE.json
configs[0] shape / SURVEY.md 8(d) config 1).  Prints ONE JSON line (see the keys below).

Timing: W >= 3 warm-up rollouts, then K rollouts, each bracketed by CUDA events on the launching
stream with a 256 MiB L2 flush between rollouts (outside the event pairs); barrier + synchronize
on both sides; max over ranks.  `value` has inputs resident in HBM; `e2e` goes through the public
Python API with pinned HOST buffers (H2D of start states and actions, D2H of the rollout stats,
inside the timed region).  `roofline` is the ensemble-forward kernel timed alone (CUDA events)
against the measured bf16 tensor peak; `cpu_baseline` is the oracle (NumPy port of the reference
CPU path; TensorFlow is not installed) on a bounded sample on this box's host cores.
"""
import argparse
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
def cpu_rollout_rate(workload, B_sample, h, reps, warmup=1):
    from oracle import ensemble as ens
    from oracle.fake_env import FakeEnv
    from oracle.static_fns import StaticFns
    from oracle.pool import SimpleReplayPool
    from oracle.rollout import rollout_model
    obs_dim, act_dim = WORKLOADS[workload]
    p, obs, acts = synthetic(workload, B_sample, h)
    env = FakeEnv(ens.OracleModel(p), StaticFns(workload), penalty_coeff=1.0, penalty_learned_var=True)
    pool = SimpleReplayPool(obs_dim, act_dim, 1250000)
    np.random.seed(0)
    times, trans = [], 0
    for i in range(warmup + reps):
        t0 = time.perf_counter()
        _, steps = rollout_model(env, pool, obs, h, lambda k, o: acts[k][:len(o)])
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt); trans = sum(steps)
    med = sorted(times)[len(times) // 2]
    return trans / med, med, trans


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.
    TensorFlow 1.14 is not installed, so this is the oracle port (kind 'port'): NumPy fp32 forward with
    BLAS-threaded matmul + the restated FakeEnv.step / ring pool / _rollout_model loop."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count()
    B_sample = args.cpu_sample
    rate, med, trans = cpu_rollout_rate(args.workload, B_sample, args.h, reps=max(args.steps, 1), warmup=max(1, min(args.warmup, 1)))
    obs_dim, act_dim = WORKLOADS[args.workload]
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': med * 1e3, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': '%s-shaped (obs %d, act %d) 7-net/5-elite 4x200 ensemble rollout, h=%d, lambda=1.0' % (args.workload, obs_dim, act_dim, args.h),
                   'rollout_batch_size': args.batch, 'rollout_length': args.h},
        'cpu_baseline': {'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': 'oracle NumPy rollout, B=%d starts x h=%d (%d transitions per timed step), median of %d' % (B_sample, args.h, trans, max(args.steps, 1))},
        'e2e': {'value': rate, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_ours(args):
    if os.environ.get('NCCL_DEBUG', 'VERSION').upper() == 'VERSION':
        os.environ['NCCL_DEBUG'] = 'WARN'          # keep stdout to the ONE json line (NCCL prints its version banner there)
    import torch
    from mopo_b200 import _lib, dist as mdist
    import __graft_entry__ as ge
    rank, world, local = mdist.init_from_env()
    if rank == 0:
        ge.build()
    mdist.barrier()
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    lib = _lib.load()
    from mopo_b200.models.constructor import construct_model
    from mopo_b200.models.fake_env import FakeEnv
    from mopo_b200.replay_pools import SimpleReplayPool
    from mopo_b200.algorithms.mopo import ModelRollout
    import mopo_b200.static as static

    obs_dim, act_dim = WORKLOADS[args.workload]
    B, h = args.batch, args.h
    prec = {'bf16': _lib.PREC_BF16, 'fp32': _lib.PREC_FP32}[args.precision]
    devnull = open(os.devnull, 'w')
    stdout, sys.stdout = sys.stdout, devnull          # the shims print like the reference; keep stdout to ONE json line
    try:
        p, obs, acts = synthetic(args.workload, B, h, seed=rank)       # per-rank starts (weak scaling), same weights below
        p0, _, _ = synthetic(args.workload, 8, h, seed=0)
        p0['mu'], p0['sigma'] = p['mu'], p['sigma']
        model = construct_model(obs_dim=obs_dim, act_dim=act_dim, hidden_dim=H_DIM, num_networks=E, num_elites=ELITES,
                                separate_mean_var=True, name='bench', device=local, precision=prec)
        if rank == 0:
            model.set_params(p0['weights'], p0['biases'], p0['var_weights'], p0['var_biases'], p0['mu'], p0['sigma'],
                             p0['max_logvar'], p0['min_logvar'])
            model._model_inds = list(range(ELITES))
        mdist.broadcast_params(model, src=0)                               # NCCL: replicate the ensemble (not timed)
        env = FakeEnv(model, static[args.workload], penalty_coeff=1.0, penalty_learned_var=True, seed=123)
        env_pool = SimpleReplayPool(obs_dim, act_dim, 16, device=local)
        pol = None
        if args.policy == 'device':                    # act = policy(obs) on the device every step
            from mopo_b200.policies import GaussianPolicy
            from oracle import policy as opol          # (random-init weights only; nothing of oracle/ is executed in the timed path)
            pol = GaussianPolicy(obs_dim, act_dim, device=local, precision=prec, seed=123)
            pol.set_params(*opol.init_params(obs_dim, act_dim, seed=0))
        ro = ModelRollout(env, env_pool, rollout_batch_size=B, rollout_schedule=(20, 100, h, h), epoch_length=1000,
                          model_train_freq=1000, model_retain_epochs=5, seed=123, policy=pol)
        ro._set_rollout_length()
        ro._reallocate_model_pool()                                        # 5 * h * B = 1.25 M rows (d4rl/base.py:19-21)
    finally:
        sys.stdout = stdout
    try:
        # probe the requested precision once
        d_obs = torch.as_tensor(obs, device=dev)
        d_acts = [torch.as_tensor(a, device=dev) for a in acts]
        ro._rollout_model(B, start_obs=d_obs, actions=None if pol else d_acts, origin_base=rank * B, quiet=True)
    except NotImplementedError as e:
        if rank == 0:
            print(json.dumps({'error': 'precision %s unavailable: %s' % (args.precision, e)}))
        return
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def timed(fn, steps, warmup, sample_clocks=False):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize(dev)
        mdist.barrier()
        clk = ClockSampler(local) if sample_clocks else None
        if clk:
            clk.start()
        evs = []
        trans = 0
        l0 = lib.mopo_launch_count()
        for _ in range(steps):
            flush.fill_(1)                                                  # L2 flush, outside the event pair
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(stream)
            trans += fn()
            e.record(stream)
            evs.append((s, e))
        torch.cuda.synchronize(dev)
        l1 = lib.mopo_launch_count()
        clocks = clk.stop() if clk else None
        mdist.barrier()
        ms = sum(s.elapsed_time(e) for s, e in evs)
        return ms, trans, l1 - l0, clocks

    def rollout_resident():
        ro._rollout_model(B, start_obs=d_obs, actions=None if pol else d_acts, origin_base=rank * B, quiet=True)
        return sum(ro._last_steps_added)

    import ctypes as C
    kt_ms, kt_n = C.c_double(0), C.c_int64(0)
    use_kt = prec == _lib.PREC_BF16
    for _ in range(args.warmup):
        rollout_resident()
    # pass 1: the headline `value` (K rollouts, CUDA events around each whole rollout, nothing recorded inside)
    ms, trans, launches, clocks = timed(rollout_resident, args.steps, 0, sample_clocks=True)
    # passes 2, 3: the same K rollouts again with CUDA events recorded on the launching stream around every launch of
    # the forward kernel (-> roofline) and of the step tail (-> what the rest of the step costs)
    tail_ms = None
    if use_kt:
        _lib.check(lib.mopo_kernel_timing(model._handle, 1))
        timed(rollout_resident, args.steps, 0)
        _lib.check(lib.mopo_kernel_time(model._handle, C.byref(kt_ms), C.byref(kt_n)))
        t_ms, t_n = C.c_double(0), C.c_int64(0)
        _lib.check(lib.mopo_kernel_timing(model._handle, 2))
        timed(rollout_resident, args.steps, 0)
        _lib.check(lib.mopo_kernel_time(model._handle, C.byref(t_ms), C.byref(t_n)))
        _lib.check(lib.mopo_kernel_timing(model._handle, 0))
        tail_ms = t_ms.value / max(t_n.value, 1)
    ms_max = mdist.max_over_ranks(ms, device=dev)
    tot = torch.tensor([float(trans)], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(tot)
    value = float(tot.item()) / (ms_max * 1e-3)

    # e2e: public API with pinned HOST buffers; H2D of starts + actions and D2H of the stats inside the timed region
    h_obs = torch.as_tensor(obs).pin_memory()
    h_acts = [torch.as_tensor(a).pin_memory() for a in acts]
    h2d = h_obs.numel() * 4 + (0 if pol else sum(a.numel() * 4 for a in h_acts))
    h_rew = torch.empty(B, 1, dtype=torch.float32).pin_memory()
    d2h = h_rew.numel() * 4

    def rollout_e2e():
        # pinned host buffers straight into the public call: it uploads the starts + step 0's actions on the compute
        # stream and the later steps' actions on a side stream, overlapped with the earlier steps
        ro._rollout_model(B, start_obs=h_obs, actions=None if pol else h_acts, origin_base=rank * B, quiet=True)
        pool = ro._model_pool                           # result read-back: the penalised rewards of the rollout's last step
        last = (pool._pointer - B) % pool._max_size
        if last + B <= pool._max_size:
            h_rew.copy_(pool.fields['rewards'][last:last + B], non_blocking=True)
        else:
            k = pool._max_size - last
            h_rew[:k].copy_(pool.fields['rewards'][last:], non_blocking=True)
            h_rew[k:].copy_(pool.fields['rewards'][:B - k], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return sum(ro._last_steps_added)

    ms_e, trans_e, _, _ = timed(rollout_e2e, args.steps, args.warmup)
    ms_e_max = mdist.max_over_ranks(ms_e, device=dev)
    tot_e = torch.tensor([float(trans_e)], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(tot_e)
    e2e_value = float(tot_e.item()) / (ms_e_max * 1e-3)

    # roofline of the dominant kernel (the tcgen05 ensemble forward): algorithmic FLOPs of the rows one launch processes
    # / its average launch duration, measured with CUDA events around every launch INSIDE the timed rollouts above
    burst, sustained, hbm, src = peaks()
    fpt = flop_per_transition(obs_dim, act_dim)
    if use_kt and kt_n.value > 0:
        ms_f = kt_ms.value / kt_n.value
        rows_per_launch = trans / (h * args.steps)
        where = 'timed inside %d rollouts of the same workload (%d launches, CUDA events on the launching stream)' % (args.steps, kt_n.value)
    else:                                               # fp32 path: one launch per layer -> time the whole forward call alone
        xn = torch.randn(B, obs_dim + act_dim, device=dev)
        mean = torch.empty(E, B, obs_dim + 1, device=dev)
        lv = torch.empty_like(mean)

        def fwd():
            _lib.check(lib.mopo_ens_forward(model._handle, xn.data_ptr(), B, mean.data_ptr(), lv.data_ptr(), prec, stream.cuda_stream))
            return B
        ms_f, _, _, _ = timed(fwd, max(args.steps, 10), 3)
        ms_f /= max(args.steps, 10)
        rows_per_launch = B
        where = 'forward call timed alone'
    achieved = fpt * rows_per_launch / (ms_f * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(ROOT, 'profiles', 'roofline_traffic.json')
    if os.path.exists(tpath) and use_kt:
        traffic = json.load(open(tpath)).get('dram_bytes_per_launch')
    roofline = {'bound': 'tensor', 'achieved': achieved, 'peak': sustained, 'unit': 'TFLOP/s', 'frac': achieved / sustained,
                'traffic': traffic,
                'kernel': 'ens_tc_kernel (tcgen05 ensemble forward + fused head, %s), %d rows/launch; %s' % (args.precision, rows_per_launch, where),
                'kernel_ms': ms_f, 'flop_per_transition': fpt,
                'peak_source': '%s bf16 sustained (kernel timed inside a long step); burst %.1f -> frac %.3f' % (src, burst, achieved / burst),
                'kernel_share_of_step': (ms_f * h * args.steps) / ms, 'tail_kernel_ms': tail_ms,
                'rollout_frac_of_sustained': value / world * fpt / 1e12 / sustained}

    if rank != 0:
        return
    cores = os.cpu_count()
    if args.no_cpu_baseline:
        cpu = None
    else:
        rate, med, tr = cpu_rollout_rate(args.workload, args.cpu_sample, h, reps=3)
        cpu = {'value': rate, 'unit': UNIT, 'cores': cores, 'kind': 'port',
               'sample': 'oracle NumPy rollout (BLAS-threaded fp32 forward + restated FakeEnv.step/pool), B=%d x h=%d = %d transitions, median of 3 (TensorFlow absent: reference arm = port)' % (args.cpu_sample, h, tr)}
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_max / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': args.precision, 'data': 'synthetic',
        'config': {'workload': '%s-shaped (obs %d, act %d) 7-net/5-elite 4x200 ensemble rollout, B=%d starts per GPU, h=%d, lambda=1.0, into a %d-row device model_pool'
                               % (args.workload, obs_dim, act_dim, B, h, ro._model_pool._max_size),
                   'rollout_batch_size': B, 'rollout_length': h, 'transitions_per_step_per_gpu': trans // args.steps,
                   'actions': 'on-device policy (2x256 ReLU Gaussian actor) every step' if pol else 'precomputed tanh(N(0,1)) arrays', 'l2': 'flushed (256 MiB write) between timed rollouts', 'parallelism': 'batch-sharded x%d, no data-path collective' % world},
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'ms_per_step': ms_e_max / args.steps},
        'gpu_launches': launches,
        'clocks': clocks,
        'roofline': roofline,
        'cpu_baseline': cpu,
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='halfcheetah', choices=sorted(WORKLOADS))
    ap.add_argument('--precision', default=os.environ.get('MOPO_BENCH_PRECISION', 'bf16'), choices=['bf16', 'fp32'])
    ap.add_argument('--batch', type=int, default=50000)
    ap.add_argument('--h', type=int, default=5)
    ap.add_argument('--cpu-sample', type=int, default=4000, help='starts in the bounded CPU-baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
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
