"""ORACLE shim: stable_baselines3 (only BaseCallback is imported, by bluesky_gym/utils/logger.py:3)."""
