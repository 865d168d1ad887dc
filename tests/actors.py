"""The reference's shipped actors (tests/golden/actors.npz, written by tests/golden/make_actors.py) as plain MLPs.

numpy for the CPU oracle, torch for the device replay. Deterministic actions as SB3's predict(deterministic=True) gives
them: SAC = action_scale * tanh(mean) (gSDE actors clip the mean to [-2, 2] first), PPO = mean clipped to the action box.
Test infrastructure: the weights are pybullet-trained policies used as a statistical pin of the rigid-body arithmetic.
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
META = json.load(open(os.path.join(HERE, 'golden', 'actors.json')))
_W = None


def weights(env):
    global _W
    if _W is None:
        _W = dict(np.load(os.path.join(HERE, 'golden', 'actors.npz')))
    return {k.split('.', 1)[1]: v for k, v in _W.items() if k.startswith(env + '.')}


def numpy_actor(env):
    w = {k: v.astype(np.float64) for k, v in weights(env).items()}
    m = META[env]
    f = np.tanh if m['activation'] == 'tanh' else (lambda x: np.maximum(x, 0))

    def act(obs):
        x = np.asarray(obs, np.float32).astype(np.float64)
        h = f(x @ w['W0'].T + w['b0']); h = f(h @ w['W1'].T + w['b1'])
        mu = h @ w['Wm'].T + w['bm']
        if m['head'] == 'clip':
            return np.clip(mu, -m['action_scale'], m['action_scale'])
        if m['head'] == 'hardtanh2_tanh':
            mu = np.clip(mu, -2.0, 2.0)
        return m['action_scale'] * np.tanh(mu)
    return act


def torch_actor(env, device):
    import torch
    w = {k: torch.tensor(v, device=device) for k, v in weights(env).items()}
    m = META[env]
    f = torch.tanh if m['activation'] == 'tanh' else torch.relu

    def act(obs):
        x = obs.to(torch.float32)
        h = f(x @ w['W0'].T + w['b0']); h = f(h @ w['W1'].T + w['b1'])
        mu = h @ w['Wm'].T + w['bm']
        if m['head'] == 'clip':
            return mu.clamp(-m['action_scale'], m['action_scale'])
        if m['head'] == 'hardtanh2_tanh':
            mu = mu.clamp(-2.0, 2.0)
        return m['action_scale'] * torch.tanh(mu)
    return act
