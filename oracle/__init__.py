"""ORACLE package: CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from here; the product (bluesky_gym_hallucination_b200) never does."""
