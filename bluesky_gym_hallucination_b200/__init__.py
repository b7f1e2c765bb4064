"""bluesky_gym_hallucination_b200 -- B200-native batched simulation core for the BlueSky-Gym
conflict-resolution envs (HorizontalCREnv-v0 / VerticalCREnv-v0 / SectorCREnv-v0 / MergeEnv-v0).
The compute path is libbsgcore.so (hand-written CUDA, sm_100a) behind the C-ABI of include/bsgcore.h;
there is no CPU fallback.  Public API (mirrors bluesky_gym/__init__.py:4-46 of the reference):

    import bluesky_gym_hallucination_b200 as bluesky_gym
    bluesky_gym.register_envs()
    env = bluesky_gym.make('HorizontalCREnv-v0', render_mode=None)      # == gymnasium.make when installed
    venv = bluesky_gym.BlueSkyVecEnv('SectorCREnv-v0', num_envs=65536, num_aircraft=50, mvp=True)
"""
__version__ = "0.1.0"

_PKG = __name__
_REGISTRATIONS = (          # id, entry point, max_episode_steps -- bluesky_gym/__init__.py:18-34
    ("HorizontalCREnv-v0", f"{_PKG}.envs:HorizontalCREnv", 300),
    ("VerticalCREnv-v0", f"{_PKG}.envs:VerticalCREnv", 300),
    ("SectorCREnv-v0", f"{_PKG}.envs:SectorCREnv", 200),
    ("MergeEnv-v0", f"{_PKG}.envs:MergeEnv", 50),
)


def register_envs():
    """Register the conflict-resolution env ids of the hot path (with gymnasium when it is installed)."""
    from .gym_compat import register
    for env_id, entry, steps in _REGISTRATIONS:
        register(id=env_id, entry_point=entry, max_episode_steps=steps)


def make(env_id, **kwargs):
    from .gym_compat import make as _make
    return _make(env_id, **kwargs)


def __getattr__(name):          # lazy: importing the package must not need torch / CUDA
    if name in ("BlueSkyVecEnv", "SB3VecEnvAdapter"):
        from . import vec_env
        return getattr(vec_env, name)
    if name == "BsgCore":
        from .core import BsgCore
        return BsgCore
    raise AttributeError(name)
