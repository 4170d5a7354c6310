"""construct_model -- same signature as the reference (mopo/models/constructor.py:7-43)."""
import numpy as np

from .fc import FC
from .bnn import BNN


def construct_model(obs_dim=11, act_dim=3, rew_dim=1, hidden_dim=200, num_networks=7,
                    num_elites=5, session=None, model_type='mlp', separate_mean_var=False,
                    name=None, load_dir=None, deterministic=False, **b200_kwargs):
    """`session` is accepted and ignored (no TF).  b200_kwargs: device=, precision=, seed=."""
    if name is None:
        name = 'BNN'
    print('[ BNN ] Name {} | Observation dim {} | Action dim: {} | Hidden dim: {}'.format(name, obs_dim, act_dim, hidden_dim))
    params = {'name': name, 'num_networks': num_networks, 'num_elites': num_elites,
              'sess': session, 'separate_mean_var': separate_mean_var, 'deterministic': deterministic}
    params.update(b200_kwargs)
    if load_dir is not None:
        print('Specified load dir', load_dir)
        params['model_dir'] = load_dir
    model = BNN(params)
    if model_type == 'identity':                                       # constructor.py:23-24
        return
    elif model_type == 'linear':
        print('[ BNN ] Training linear model')                          # constructor.py:25-27
        model.add(FC(obs_dim + rew_dim, input_dim=obs_dim + act_dim, weight_decay=0.000025))
    elif model_type == 'mlp':                                          # constructor.py:28-36
        print('[ BNN ] Training non-linear model | Obs: {} | Act: {} | Rew: {}'.format(obs_dim, act_dim, rew_dim))
        model.add(FC(hidden_dim, input_dim=obs_dim + act_dim, activation="swish", weight_decay=0.000025))
        model.add(FC(hidden_dim, activation="swish", weight_decay=0.00005))
        model.add(FC(hidden_dim, activation="swish", weight_decay=0.000075))
        model.add(FC(hidden_dim, activation="swish", weight_decay=0.000075))
        model.add(FC(obs_dim + rew_dim, weight_decay=0.0001))
        if separate_mean_var:
            model.add(FC(obs_dim + rew_dim, input_dim=hidden_dim, weight_decay=0.0001), var_layer=True)
    else:
        raise ValueError('unknown model_type %r' % (model_type,))
    if load_dir is not None:
        model.model_loaded = True                                       # constructor.py:38-39
    model.finalize(None, {"learning_rate": 0.001}, obs_dim=obs_dim)
    print('[ BNN ] Model: {}'.format(model))
    return model


def format_samples_for_training(samples):
    """constructor.py:46-57: (obs|act) -> (rew|delta_obs)."""
    obs, act, next_obs, rew = samples['observations'], samples['actions'], samples['next_observations'], samples['rewards']
    delta_obs = next_obs - obs
    if hasattr(obs, 'is_cuda'):                  # device pool fields stay in HBM (torch tensors)
        import torch
        return torch.cat((obs, act), dim=-1), torch.cat((rew, delta_obs), dim=-1)
    inputs = np.concatenate((obs, act), axis=-1)
    outputs = np.concatenate((rew, delta_obs), axis=-1)
    return inputs, outputs
