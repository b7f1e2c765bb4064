"""ORACLE (test infrastructure, not product code): the abstract surface of stable_baselines3.common.vec_env.VecEnv
(stable-baselines3 2.x, un-vendored dependency of the reference's scripts/multi_processing_example.py:38-54), restated
from its published interface: names and signatures only, so that tests can check that SB3VecEnvAdapter would survive
`isinstance(env, VecEnv)` users (VecMonitor, VecNormalize, make_vec_env).  Nothing here computes anything."""
import abc


class VecEnv(abc.ABC):
    def __init__(self, num_envs, observation_space, action_space):
        self.num_envs, self.observation_space, self.action_space = num_envs, observation_space, action_space
        self.render_mode = None

    @abc.abstractmethod
    def reset(self): ...

    @abc.abstractmethod
    def step_async(self, actions): ...

    @abc.abstractmethod
    def step_wait(self): ...

    @abc.abstractmethod
    def close(self): ...

    @abc.abstractmethod
    def get_attr(self, attr_name, indices=None): ...

    @abc.abstractmethod
    def set_attr(self, attr_name, value, indices=None): ...

    @abc.abstractmethod
    def env_method(self, method_name, *method_args, indices=None, **method_kwargs): ...

    @abc.abstractmethod
    def env_is_wrapped(self, wrapper_class, indices=None): ...

    def step(self, actions):
        self.step_async(actions)
        return self.step_wait()

    def seed(self, seed=None): ...

    def get_images(self): ...

    def render(self, mode=None): ...
