"""bluesky_gym_hallucination_b200 -- B200-native batched simulation core for the BlueSky-Gym
conflict-resolution envs (HorizontalCREnv-v0 / VerticalCREnv-v0 / SectorCREnv-v0).
The compute path is libbsgcore.so (hand-written CUDA, sm_100a) behind the C-ABI of include/bsgcore.h."""
__version__ = "0.1.0"
