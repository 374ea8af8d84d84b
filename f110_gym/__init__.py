"""Drop-in replacement for the reference's `f110_gym` package (gym/f110_gym/__init__.py:1-5):
registers 'f110-v0' -> f110_gym.envs:F110Env when gym/gymnasium is installed, and offers
`f110_gym.make` for environments without gym."""
try:   # reference: gym==0.19
    from gym.envs.registration import register
    register(id='f110-v0', entry_point='f110_gym.envs:F110Env')
except Exception:   # gym not installed (as on the build/GPU boxes)
    pass


def make(env_id='f110_gym:f110-v0', **kwargs):
    """gym.make('f110_gym:f110-v0', **kwargs) without gym."""
    if env_id.split(':')[-1] != 'f110-v0':
        raise ValueError('unknown environment id %s' % env_id)
    from f110_gym.envs import F110Env
    return F110Env(**kwargs)
