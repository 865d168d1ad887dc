"""Host-side mirror of the reference's Gymnasium interface (urgym_b200/envs.py), CPU-checkable parts."""
import numpy as np
import pytest

from urgym_b200 import envs as E
from urgym_b200._lib import ENV_TYPES


def test_specs_cover_exactly_the_in_scope_envs():
    assert {v[0] for v in E.SPECS.values()} == set(ENV_TYPES)
    assert set(E._CLASSES) == set(E.SPECS)
    for env_id, cls in E._CLASSES.items():
        assert cls.env_id == env_id and cls.__name__ == E.SPECS[env_id][0]


def test_spaces_shapes_and_dtypes(built):
    from urgym_b200 import _lib
    L = _lib.lib()
    for env_id, (name, _, _, _) in E.SPECS.items():
        o, a = E._spaces(env_id)
        assert o.dtype == np.float64 and a.dtype == np.float32
        assert o.shape == (L.urgym_obs_dim(ENV_TYPES[name]),) and a.shape == (L.urgym_act_dim(ENV_TYPES[name]),)
        assert np.all(o.low <= o.high) and np.all(a.low <= a.high)
        s = a.sample()
        assert s.dtype == np.float32 and a.contains(s)


def test_out_of_scope_ids_say_so():
    for env_id in E.UNSUPPORTED:
        with pytest.raises(NotImplementedError):
            E.make(env_id)
    with pytest.raises(KeyError):
        E.make('URGym/Nope-v0')


def test_no_cpu_fallback_for_the_env_classes():
    import torch
    if torch.cuda.is_available():
        pytest.skip('checks the no-device behaviour')
    from urgym_b200._lib import UrgymError
    with pytest.raises(UrgymError):
        E.make('URGym/BallBalance-v0', render_mode='train')
    with pytest.raises(UrgymError):
        E.URGymVectorEnv('URGym/Golf-v0', 8)
    with pytest.raises(ValueError):
        E.URGymVectorEnv('URGym/Nope-v0', 8)
    with pytest.raises(NotImplementedError):
        E.make('URGym/Box-v0', button_touch_mode='fingers')   # only the registered default 'any' is backed
    from urgym_b200.sb3_vec_env import URGymSB3VecEnv
    with pytest.raises(UrgymError):
        URGymSB3VecEnv('URGym/CubesPush-v0', 8)
