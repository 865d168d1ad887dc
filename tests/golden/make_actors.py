#!/usr/bin/env python
"""Extracts the actor networks of the reference's shipped SB3 checkpoints into tests/golden/actors.npz.

Run where /root/reference exists (it does not on the GPU box, so the weights travel as this fixture):

    python tests/golden/make_actors.py

Sources (binary checkpoints, not source code): /root/reference/4000000.zip (PPO, TwoBallsBalance, 9 -> 64 -> 64 -> 3, tanh),
/root/reference/policies_final/{cubes_push_v0,golf_v0,golf_joints_v0}/SAC_policy.zip (SAC actors, obs -> 256 -> 256 -> act,
ReLU, tanh-squashed mean; the CubesPush actor was trained with gSDE, whose mean head ends in Hardtanh(-2, 2)). Only
`policy.pth` is read, with torch.load(weights_only=True); nothing is unpickled. The Monitor statistics of the last 100
training episodes on pybullet go into actors.json as the targets of the replay.
"""
import io
import json
import os
import zipfile

import numpy as np
import torch

REF = os.environ.get('URGYM_REFERENCE', '/root/reference')
HERE = os.path.dirname(os.path.abspath(__file__))
CASES = {
    'TwoBallsBalance': ('4000000.zip', 'ppo'),
    'CubesPush': ('policies_final/cubes_push_v0/SAC_policy.zip', 'sac'),
    'Golf': ('policies_final/golf_v0/SAC_policy.zip', 'sac'),
    'GolfJoints': ('policies_final/golf_joints_v0/SAC_policy.zip', 'sac'),
}
# BASELINE.md section 2 (ep_info_buffer / ep_success_buffer of each checkpoint, last 100 training episodes on pybullet)
TRAINED = {
    'TwoBallsBalance': dict(mean_return=64.07, mean_length=18.09, success_rate=None,
                            note='trained on an older env revision (ball radius 0.015, now 0.003): indicative only'),
    'CubesPush': dict(mean_return=9.67, mean_length=8.13, success_rate=1.00),
    'Golf': dict(mean_return=9.01, mean_length=13.18, success_rate=0.99),
    'GolfJoints': dict(mean_return=10.19, mean_length=17.39, success_rate=0.98),
}


def main():
    out, meta = {}, {}
    for env, (rel, algo) in CASES.items():
        z = zipfile.ZipFile(os.path.join(REF, rel))
        sd = torch.load(io.BytesIO(z.read('policy.pth')), map_location='cpu', weights_only=True)
        g = lambda k: sd[k].float().numpy()
        if algo == 'ppo':
            layers = [('mlp_extractor.policy_net.0', 'W0', 'b0'), ('mlp_extractor.policy_net.2', 'W1', 'b1'), ('action_net', 'Wm', 'bm')]
            kind = dict(activation='tanh', head='clip')          # PPO: action = mean, clipped to the action box
        else:
            sde = 'actor.mu.0.weight' in sd
            layers = [('actor.latent_pi.0', 'W0', 'b0'), ('actor.latent_pi.2', 'W1', 'b1'), ('actor.mu.0' if sde else 'actor.mu', 'Wm', 'bm')]
            kind = dict(activation='relu', head='hardtanh2_tanh' if sde else 'tanh')   # SAC: action = 0.1 * tanh(mean)
        for src, wn, bn in layers:
            out['%s.%s' % (env, wn)] = g(src + '.weight'); out['%s.%s' % (env, bn)] = g(src + '.bias')
        meta[env] = dict(checkpoint=rel, algo=algo, obs_dim=int(out[env + '.W0'].shape[1]), act_dim=int(out[env + '.Wm'].shape[0]),
                         action_scale=0.1, pybullet_training_last_100=TRAINED[env], **kind)
    np.savez_compressed(os.path.join(HERE, 'actors.npz'), **out)
    with open(os.path.join(HERE, 'actors.json'), 'w') as f:
        json.dump(meta, f, indent=1)
    print({k: v.shape for k, v in out.items()})


if __name__ == '__main__':
    main()
