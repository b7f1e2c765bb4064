"""Samples of the UNMODIFIED reference generators (the three envs' reset()), for the distributional test of the device
reset kernels (tests/test_gpu_vecenv.py::test_reset_distributions_match_the_reference).  Run HERE (needs /root/reference):

    python tests/golden/make_reset_samples.py

For every env kind, M resets of the reference class (on top of oracle/refshim) are recorded as the per-aircraft arrays
lat / lon / alt / hdg / tas (float32, padded to n slots), the live count and the env-level slots the generator draws
(waypoint, target altitude, sector polygon) -> tests/golden/reset_samples_{H,V,S}.npz.
"""
import numpy as np
import make_golden as mg           # sets up sys.path (refshim + /root/reference), snapshot()

bs, gym, orc = mg.bs, mg.gym, mg.orc


def sample(kind, M, n, seed):
    np.random.seed(seed)
    env = gym.make(mg.ENV_ID[kind], render_mode=None)
    F = orc.F
    keep = ['lat', 'lon', 'alt', 'hdg', 'tas']
    ac = np.zeros((M, len(keep), n), np.float64); nac = np.zeros(M, np.int32); ev = np.zeros((M, orc.NE))
    for m in range(M):
        env.reset()
        a, _, nl, e = mg.snapshot(kind, env, n)
        ac[m] = a[[F[k] for k in keep]]; nac[m] = nl; ev[m] = e
    return dict(ac=ac, nac=nac, ev=ev, fields=np.array(keep))


def main():
    mg.bluesky_gym.register_envs()
    for kind, M, n, seed in (('H', 600, 6, 501), ('V', 600, 6, 502), ('S', 300, 32, 503)):
        R = sample(kind, M, n, seed)
        np.savez_compressed(mg.os.path.join(mg.HERE, f'reset_samples_{kind}.npz'), **R)
        print(kind, R['ac'].shape, 'nac', R['nac'].min(), R['nac'].mean(), R['nac'].max())


if __name__ == '__main__':
    main()
