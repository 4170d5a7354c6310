"""BASELINE.json config 5 on one GPU: the MOPO epoch loop's device work -- one model rollout (B starts x h steps, actor on
the device, into the model pool) + n SAC updates (batch 256 = 12 env + 244 model rows drawn on the device) + the actor
export back into the rollout policy -- timed with CUDA events; beside it the float64 oracle port of ONE SAC update on the
host cores.  usage: epoch_probe.py [B] [h] [updates] [epochs]

Under torchrun (N ranks, one per GPU) the rollout's B starts are sharded over the ranks, every rank all-gathers the new
model-pool rows of all shards over NCCL (the pool gather BASELINE.json names) into its copy of the model pool, and the SAC
updates run replicated with the same seeds -- identical actors on every rank, so no parameter exchange is needed; the
time is the max over ranks."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mopo_b200 import _lib                                            # noqa: E402
import mopo_b200.static as static                                     # noqa: E402
from mopo_b200.models.constructor import construct_model             # noqa: E402
from mopo_b200.models.fake_env import FakeEnv                        # noqa: E402
from mopo_b200.replay_pools import SimpleReplayPool                  # noqa: E402
from mopo_b200.algorithms.mopo import ModelRollout                   # noqa: E402
from mopo_b200.algorithms.sac import SACUpdate                       # noqa: E402
from mopo_b200.policies import GaussianPolicy                        # noqa: E402
from mopo_b200 import dist as mdist                                  # noqa: E402
from oracle import ensemble as ens, sac as osac                      # noqa: E402


def main():
    B, h, n_upd, epochs = [int(a) for a in sys.argv[1:5]] + [50000, 5, 1000, 3][len(sys.argv) - 1:]
    od, ad = 17, 6
    rank, world, local = mdist.init_from_env()
    torch.cuda.set_device(local)
    dev = local
    lo, hi = mdist.shard_range(B, rank, world)
    Bl = hi - lo
    p, obs, _ = ens.synthetic_case('halfcheetah', 100000, seed=0)
    rng = np.random.RandomState(1)
    env_pool = SimpleReplayPool(od, ad, 100000, device=dev)
    env_pool.add_samples({'observations': obs, 'actions': np.tanh(rng.standard_normal((100000, ad))).astype(np.float32),
                          'next_observations': obs + 0.1 * rng.standard_normal(obs.shape).astype(np.float32),
                          'rewards': rng.standard_normal((100000, 1)).astype(np.float32), 'terminals': np.zeros((100000, 1), bool)})
    model = construct_model(obs_dim=od, act_dim=ad, separate_mean_var=True, device=dev, precision=_lib.PREC_BF16)
    model.set_params(p['weights'], p['biases'], p['var_weights'], p['var_biases'], p['mu'], p['sigma'], p['max_logvar'], p['min_logvar'])
    model._model_inds = [0, 1, 2, 3, 4]
    env = FakeEnv(model, static['halfcheetah'], penalty_coeff=1.0, penalty_learned_var=True, seed=3)
    pol = GaussianPolicy(od, ad, device=dev, precision=_lib.PREC_BF16, seed=5)
    sac = SACUpdate(od, ad, device=dev, seed=7)
    sac.export_policy(pol)
    ro = ModelRollout(env, env_pool, rollout_batch_size=Bl, rollout_schedule=(0, 1, h, h), epoch_length=1000, model_train_freq=1000,
                      model_retain_epochs=5 if world == 1 else 1, policy=pol, seed=9)
    ro._set_rollout_length()
    ro._reallocate_model_pool()
    shard_pool = ro._model_pool                       # this rank's rows
    full_pool = shard_pool if world == 1 else SimpleReplayPool(od, ad, 5 * B * h, device=dev)
    fields = ('observations', 'actions', 'next_observations', 'rewards', 'terminals')

    def epoch():
        ro._rollout_model(Bl, quiet=True, origin_base=lo)
        if world > 1:                                 # NCCL: every rank receives every shard's new rows (equal counts: halfcheetah never terminates)
            n_new = Bl * h
            ptr = shard_pool._pointer
            rows = (torch.arange(ptr - n_new, ptr, device=shard_pool.device) % shard_pool._max_size)
            got = mdist.allgather_rows({k: shard_pool.fields[k][rows] for k in fields})
            full_pool.add_samples(got)
        sac.train_from_pools(env_pool, full_pool, n_updates=n_upd)
        sac.export_policy(pol)

    l0 = _lib.load().mopo_launch_count()
    epoch()
    torch.cuda.synchronize()
    launches = _lib.load().mopo_launch_count() - l0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    mdist.barrier()
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(epochs):
        epoch()
    ev[1].record()
    torch.cuda.synchronize()
    ev[2].record()
    sac.train_from_pools(env_pool, full_pool, n_updates=n_upd)
    ev[3].record()
    torch.cuda.synchronize()
    mdist.barrier()
    ms_epoch = mdist.max_over_ranks(ev[0].elapsed_time(ev[1]) / epochs, device=torch.device('cuda', dev))
    ms_upd = ev[2].elapsed_time(ev[3]) / n_upd
    # host baseline: the float64 NumPy port of one SAC update
    st = osac.init_state(od, ad, seed=0)
    opt = osac.adam_init(st)
    hp = dict(osac.DEFAULT_HP, target_entropy=-float(ad))
    batch = {'observations': obs[:256].astype(np.float64), 'actions': np.tanh(rng.standard_normal((256, ad))),
             'next_observations': obs[256:512].astype(np.float64), 'rewards': rng.standard_normal((256, 1)), 'terminals': np.zeros((256, 1), bool)}
    k = 5
    t0 = time.perf_counter()
    for _ in range(k):
        osac.update(st, opt, batch, rng.standard_normal((256, ad)), rng.standard_normal((256, ad)), hp)
    cpu_ms = (time.perf_counter() - t0) / k * 1e3
    check = sac.get_params()['pi']['W'][0]
    if world > 1:                                     # replicated SAC: every rank must hold the same actor
        t = torch.as_tensor(check).to(torch.device('cuda', dev)); ref = t.clone()
        torch.distributed.broadcast(ref, src=0)
        assert torch.equal(t, ref), 'actors diverged across ranks'
    if rank != 0:
        mdist.shutdown()
        return
    print(json.dumps({'n_gpus': world, 'workload': 'mopo epoch loop (config 5, %d GPU): rollout B=%d h=%d with the device actor + %d SAC updates (batch 256 = 12 env + 244 model) + actor export' % (world, B, h, n_upd),
                      'ms_per_epoch': ms_epoch, 'ms_per_sac_update': ms_upd, 'sac_updates_per_s': 1e3 / ms_upd,
                      'rollout_transitions_per_epoch': B * h, 'gpu_launches_per_epoch': launches,
                      'pool_rows': full_pool.size, 'cpu_port_ms_per_sac_update': cpu_ms, 'cpu_cores': os.cpu_count()}))
    mdist.shutdown()


if __name__ == '__main__':
    main()
