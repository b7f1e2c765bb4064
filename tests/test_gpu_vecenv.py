"""GPU tests of the host-side mirror of the reference interface: the registered single-env ids, the batched
VectorEnv (next-step auto-reset) and the SB3-style adapter (same-step auto-reset, terminal_observation)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_registered_single_envs_speak_the_reference_api():
    import bluesky_gym_hallucination_b200 as bluesky_gym
    bluesky_gym.register_envs()
    for env_id, act_dim, third in (('HorizontalCREnv-v0', 1, 'average_drift'), ('VerticalCREnv-v0', 1, 'final_altitude'),
                                   ('SectorCREnv-v0', 2, 'average_drift')):
        env = bluesky_gym.make(env_id, render_mode=None)
        obs, info = env.reset(seed=3)
        assert set(obs) == set(env.observation_space.spaces.keys())
        for k, v in obs.items():
            assert v.dtype == np.float64 and v.shape == env.observation_space[k].shape
        assert set(info) == {'total_reward', 'total_intrusions', third}
        if third == 'average_drift':
            assert np.isnan(info[third])                       # mean of an empty list right after reset
        tot = 0.0
        for _ in range(5):
            obs, rew, term, trunc, info = env.step(np.zeros(act_dim))
            assert isinstance(rew, float) and term in (0, 1) and isinstance(trunc, bool)
            tot += rew
        assert abs(info['total_reward'] - tot) < 1e-9
        env.close()
    with pytest.raises(NotImplementedError):
        bluesky_gym.make('HorizontalCREnv-v0', render_mode='human')


def test_horizontal_straight_flight_reaches_the_waypoint_like_the_reference_logs():
    """Anchor from the shipped logs: flying straight, Horizontal episodes end after ~24-37 steps."""
    import bluesky_gym_hallucination_b200 as bluesky_gym
    bluesky_gym.register_envs()
    env = bluesky_gym.make('HorizontalCREnv-v0')
    lens = []
    for ep in range(4):
        env.reset(seed=100 + ep)
        for n in range(1, 301):
            _, _, term, trunc, _ = env.step(np.zeros(1))
            if term or trunc:
                break
        lens.append(n)
    assert all(22 <= n <= 40 for n in lens), lens
    env.close()


def test_time_limit_truncates_single_env():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    v = BlueSkyVecEnv('SectorCREnv-v0', 8, max_episode_steps=3)
    v.reset(seed=1)
    for s in range(3):
        obs, rew, term, trunc, info = v.step(np.zeros((8, 2)))
    assert trunc.all()
    v.close()


def test_vector_env_next_step_autoreset():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv
    E = 64
    v = BlueSkyVecEnv('HorizontalCREnv-v0', E, num_aircraft=6, precision='fp32', seed=5)
    obs, info = v.reset(seed=5)
    assert obs['intruder_distance'].shape == (E, 5) and obs['cos_drift'].shape == (E, 1)
    done_prev = np.zeros(E, bool)
    ended = 0
    for s in range(60):
        obs, rew, term, trunc, info = v.step(np.zeros((E, 1)))
        # the step after a done returns the first observation of the new episode: reward 0, flags 0, counters 0
        assert np.all(rew[done_prev] == 0.0) and not term[done_prev].any() and not trunc[done_prev].any()
        assert np.all(info['total_reward'][done_prev] == 0.0)
        assert np.all(obs['waypoint_distance'][done_prev, 0] > 95.0 / 150.0)      # fresh waypoint 100-149 km away
        done_prev = term | trunc
        ended += int(done_prev.sum())
    assert ended >= E            # straight flight ends every episode within ~37 steps
    t = v.step_tensors(torch.zeros((E, 1), dtype=torch.float32, device='cuda'))
    assert t[0].is_cuda and t[0].shape == (E, 28)
    v.close()


def test_sb3_adapter_same_step_autoreset():
    from bluesky_gym_hallucination_b200.vec_env import BlueSkyVecEnv, SB3VecEnvAdapter
    E = 32
    a = SB3VecEnvAdapter(BlueSkyVecEnv('HorizontalCREnv-v0', E, precision='fp32', seed=9))
    obs = a.reset()
    seen = 0
    for s in range(45):
        obs, rew, dones, infos = a.step(np.zeros((E, 1)))
        assert set(infos[0]) >= {'total_reward', 'total_intrusions', 'average_drift', 'TimeLimit.truncated'}
        for i in np.where(dones)[0]:
            seen += 1
            assert 'terminal_observation' in infos[i]
            assert infos[i]['terminal_observation']['waypoint_distance'][0] < 5.0 / 150.0    # reached the waypoint
            assert obs['waypoint_distance'][i, 0] > 95.0 / 150.0                             # already the new episode
    assert seen >= E // 2
    a.close()
